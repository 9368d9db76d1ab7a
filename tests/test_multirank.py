"""CPU, world_size 2 over gloo: the N>1 path of bench.py -- contiguous sharding by instance index, counter-based inputs,
MAX-over-ranks timing reduction and the one optional gather of results.  No collective sits on the hot path."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, n_total, q):
    sys.path.insert(0, ROOT)
    from quadrupedwholebodycontroller_b200 import workloads as wl
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    first, count = wl.shard_range(n_total, rank, world)
    st, ref = wl.randomised(count, seed=1234, first=first, contacts="mixed")
    # what bench.py does with its timings: MAX over ranks
    t = torch.tensor([1.0 + rank], dtype=torch.float64); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    # optional gather of per-robot results (here: a checksum of the inputs standing in for tau)
    local = torch.from_numpy(st["q"].sum(axis=0) + ref["contacts"].sum(axis=0))
    sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([count]))
    bufs = [torch.zeros(int(s), dtype=torch.float64) for s in sizes]
    dist.all_gather(bufs, local) if len(set(int(s) for s in sizes)) == 1 else [dist.broadcast(bufs[r] if r != rank else local, r) for r in range(world)]
    if len(set(int(s) for s in sizes)) != 1:
        bufs[rank] = local
    if rank == 0:
        q.put((float(t), torch.cat(bufs).numpy(), [int(s) for s in sizes]))
    dist.barrier(); dist.destroy_process_group()


def _run(n_total):
    ctx = mp.get_context("spawn"); q = ctx.Queue(); port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, n_total, q)) for r in range(2)]
    [p.start() for p in ps]
    tmax, gathered, sizes = q.get(timeout=120)
    [p.join(60) for p in ps]
    assert all(p.exitcode == 0 for p in ps)
    return tmax, gathered, sizes


def test_two_ranks_cover_the_batch_exactly_once(wl):
    n_total = 37
    tmax, gathered, sizes = _run(n_total)
    st, ref = wl.randomised(n_total, seed=1234, contacts="mixed")
    assert sizes == [19, 18] and tmax == 2.0
    assert np.array_equal(gathered, st["q"].sum(axis=0) + ref["contacts"].sum(axis=0))


def test_shard_ranges_partition(wl):
    for n, w in ((1 << 20, 8), (4096, 3), (5, 8)):
        r = [wl.shard_range(n, k, w) for k in range(w)]
        assert r[0][0] == 0 and sum(c for _, c in r) == n
        assert all(r[k][0] + r[k][1] == r[k + 1][0] for k in range(w - 1))
