"""CPU: the C-ABI library builds for sm_100a, loads, exports every symbol include/wbc_b200.h declares, and refuses to run
without a GPU (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "wbc_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(wbc_[a-z_0-9]+)\s*\(", src)))


def test_header_and_binding_agree(wbc):
    assert _declared() == sorted(wbc.EXPORTS)


def test_library_exports_every_declared_symbol(wbc):
    lib = wbc.load_library()
    for name in _declared():
        assert hasattr(lib, name), name


def test_only_sm100a_code_is_embedded(wbc):
    import subprocess
    from quadrupedwholebodycontroller_b200.build import LIB_PATH
    out = subprocess.run(["cuobjdump", "-lelf", LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out and not re.search(r"sm_(?!100a)\d+", out), out


def test_no_cpu_fallback(wbc):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        wbc.WholeBodyController(4)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "quadrupedwholebodycontroller_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "import oracle" not in txt and "from oracle" not in txt and "liboracle" not in txt and "qpOASES.hpp" not in txt, f


def test_default_params_match_reference_yaml(wbc):
    p = wbc.default_params()      # config/params_controller.yaml:1-11
    assert (p.friction, p.loopRate, p.maxTorque, p.kpValue, p.kpValueZ, p.kdValue, p.kiValue, p.kpSwingValue, p.kdSwingValue, p.slackWeight) == \
           (1.0, 400.0, 80.0, 6000.0, 10000.0, 1800.0, 0.0, 250.0, 20.0, 1000.0)
    assert list(p.initialReferencePose) == [0.0, 0.0, 0.5, 0.0, 0.0, 0.0]
