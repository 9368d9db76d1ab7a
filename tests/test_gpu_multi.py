"""GPU (-m gpu): the library-level multi-GPU object (wbc_multi_*).  With one visible GPU the same device is listed twice --
the sharding, the strided slice copies and the gather are exercised all the same; with two or more GPUs the shards really sit on different devices
and the gather goes into device memory of GPU 0 (peer copies)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [97, 256])
def test_sharded_over_devices_is_bit_identical_to_one_device(wbc, wl, n):
    import torch
    ndev = torch.cuda.device_count()
    devices = [0, 1] if ndev >= 2 else [0, 0]
    if n == 256 and ndev >= 4:
        devices = [0, 1, 2, 3]
    st, ref = wl.randomised(n, seed=21, contacts="mixed", switching_frac=0.1)
    one = wbc.WholeBodyController(n); multi = wbc.MultiGpuController(n, devices)
    sh = multi.shards()
    assert [s[1] for s in sh] == [wl.shard_range(n, k, len(devices))[0] for k in range(len(devices))]
    assert [s[2] for s in sh] == [wl.shard_range(n, k, len(devices))[1] for k in range(len(devices))] and sum(s[2] for s in sh) == n
    s = st
    for t in range(3):
        if t:
            s = wl.advance_state(s)
        if t == 2:
            ref = dict(ref); ref["contacts"] = np.ascontiguousarray(np.roll(ref["contacts"], 1, axis=0)); ref["switching"] = None      # derived per shard like wbc.cpp:176-184
        one.setJointAndBaseState(**s); one.setReference(**ref); one.step()
        multi.setJointAndBaseState(**s); multi.setReference(**ref); multi.step()
        a, b = one.getOutputs(), multi.getOutputs()
        for k in a:
            assert np.array_equal(a[k], b[k]), (t, k)
    # the optional gather: results of all shards into device memory of GPU 0
    tau = torch.zeros(12, n, dtype=torch.float64, device="cuda:0"); rc = torch.zeros(n, dtype=torch.int32, device="cuda:0")
    multi.getOutputs(tau=tau, rc=rc)
    assert np.array_equal(tau.cpu().numpy(), a["tau"]) and np.array_equal(rc.cpu().numpy(), a["rc"])
    assert multi.lastGatherMs() >= 0.0
    one.close(); multi.close()
