"""CPU, world_size 2, gloo: the host-side logic of the multi-GPU path -- track sharding and the one
collective (all-reduce SUM of the per-shard NLL grid vector followed by a common argmin).  The
per-track NLL is computed with the oracle here (no GPU in this suite); on GPUs the same code path
feeds libeksb's NLL sums into the same all-reduce (tests/test_gpu_core.py checks those sums)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from eks_b200 import selection as sel
from oracle import restated as o


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _problem(B=7, T=40, p=4, n=3):
    rng = np.random.default_rng(0)
    C = np.linalg.qr(rng.normal(size=(B, p, n)))[0]
    y = rng.normal(size=(B, T, p)); ev = np.exp(rng.normal(size=(B, T, p)))
    return dict(C=C, y=y, ev=ev, A=np.eye(n), Q=0.05 * np.eye(n), S0=np.eye(n), m0=np.zeros(n), grid=[0.1, 1.0, 10.0])


def _local_sums(pb, lo, hi):
    p = pb["y"].shape[2]
    out = np.zeros(len(pb["grid"]))
    for gi, s in enumerate(pb["grid"]):
        for b in range(lo, hi):
            out[gi] += o.nll(pb["y"][b], pb["m0"], pb["S0"], pb["C"][b], np.eye(p), pb["A"], s * pb["Q"], pb["ev"][b])
    return out


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pb = _problem()
    lo, hi = sel.shard_bounds(pb["y"].shape[0], rank, world)
    vec = torch.as_tensor(_local_sums(pb, lo, hi))
    sel.allreduce_sum_(vec)
    q.put((rank, lo, hi, vec.numpy().copy()))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_bounds_cover_everything():
    for B in (1, 7, 64, 16384):
        for world in (1, 2, 3, 8):
            spans = [sel.shard_bounds(B, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == B
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1


@pytest.mark.timeout(120)
def test_nll_allreduce_world2_gloo():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    [p.start() for p in procs]
    res = sorted(q.get(timeout=100) for _ in range(world))
    [p.join(30) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    pb = _problem()
    want = _local_sums(pb, 0, pb["y"].shape[0])
    for rank, lo, hi, vec in res:
        np.testing.assert_allclose(vec, want, rtol=1e-12)      # every rank holds the global sums
    assert res[0][1:3] == (0, 4) and res[1][1:3] == (4, 7)
    assert int(np.argmin(res[0][3])) == int(np.argmin(res[1][3]))
