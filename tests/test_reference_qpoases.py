"""Pins the qpOASES half of the oracle: the reference's OWN unit tests (lib/qpOASES/testing/cpp, the ones SURVEY.md section 8c lists for
SQProblem and its bookkeeping) are compiled from the sources where they lie under /root/reference, linked against oracle/_ref/libqpoases_ref.a
-- the very library the oracle drives -- and must all pass (exit code 0 = TEST_PASSED, UnitTesting.hpp).  Needs the reference mount (build
container); skipped where it does not exist."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/lib/qpOASES/testing/cpp"
TESTS = ["test_example1", "test_example1a", "test_example2", "test_identitySqproblem", "test_janick1", "test_janick2", "test_indexlist", "test_matrices",
         "test_gradientShift"]


@pytest.fixture(scope="module")
def built(orc):
    if not os.path.isdir(REF):
        pytest.skip("reference mount not available: qpOASES' own tests are built from it")
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "qptests", "CXX=g++"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return os.path.join(ROOT, "oracle", "_ref", "qptests")


@pytest.mark.parametrize("name", TESTS)
def test_reference_unit_test_passes(built, name):
    r = subprocess.run([os.path.join(built, name)], cwd=REF, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, (name, r.stdout[-500:], r.stderr[-500:])
