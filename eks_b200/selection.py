"""Smoothing-parameter selection by negative log-likelihood, z-scores, and the multi-GPU plumbing
for them.  The reference has none of this (its smoothing parameter is a CLI
flag, scripts/multicam_example.py:38-43); definitions are the build's own:

    NLL(s) = 1/2 sum_t [ log det Sig_t + nu_t^T Sig_t^-1 nu_t + p log 2 pi ],  Sig_t = R_t + C S_pred C^T
    (the matrix of eks/ensemble_kalman.py:112 and the innovation of :102), with Q = s * Q0.

Tracks shard independently across GPUs (one process per GPU); the ONLY collective is one
all-reduce(SUM) of the (n_grid,) float64 vector of per-GPU NLL sums, after which every rank takes
the same argmin.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

from . import _lib, device as dev
from ._lib import check, lib


def shard_bounds(n_tracks, rank, world):
    """Contiguous block of tracks owned by `rank` (sizes differ by at most one)."""
    base, extra = divmod(n_tracks, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def allreduce_sum_(vec):
    """In-place SUM over all ranks when torch.distributed is initialised (NCCL on GPUs, gloo in CPU tests)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(vec, op=dist.ReduceOp.SUM)
    return vec


def nll_grid(obs, prm, B, T, s_grid):
    """(n_grid,B) float64 NLL of every local track for Q = s*prm.Q, one launch (grid.y = n_grid)."""
    q = torch.as_tensor(np.asarray(s_grid, dtype=np.float64), dtype=obs.dtype).to(obs.device)
    nll, status = dev.kf_nll(obs, prm, B, T, q_scale=q)     # time-split when the batch has few tiles and long tracks
    return nll, status


def nll_sums(nll):
    """(n_grid,B) -> (n_grid,) on the device with the library's deterministic row reduction."""
    G, B = nll.shape
    out = torch.empty(G, dtype=torch.float64, device=nll.device)
    check(lib().eksb_reduce_rows_f64(nll.data_ptr(), G, B, out.data_ptr(),
                                     torch.cuda.current_stream(nll.device).cuda_stream), "eksb_reduce_rows")
    return out


def select_smooth_param(obs, prm, B, T, s_grid):
    """Grid search: local NLL sums -> all-reduce over track shards -> argmin.
    Returns (best_s, total_nll (n_grid,) host array)."""
    nll, _ = nll_grid(obs, prm, B, T, s_grid)
    tot = allreduce_sum_(nll_sums(nll))
    tot_h = tot.cpu().numpy()
    return float(np.asarray(s_grid)[int(np.argmin(tot_h))]), tot_h


def refine_smooth_param(obs, prm, B, T, lo, hi, iters=24, candidates=2):
    """Golden-section search of the summed NLL over log s in [lo, hi] (the same objective and the same one
    all-reduce per evaluation as select_smooth_param; every rank takes the same steps).  Each step evaluates
    `candidates` = 2 interior points in one launch, through eksb_kf_nll_ws with ONE workspace kept for the whole
    search: for a few keypoints x a long video an evaluation is time-split (include/eksb.h) -- 1.3 ms instead of
    20.7 ms for one keypoint x 100 000 frames.  Returns (best_s, evaluations [(s, nll)], ...)."""
    assert candidates == 2
    phi = (np.sqrt(5.0) - 1.0) / 2.0
    a, b = np.log(lo), np.log(hi)
    ws = dev.nll_workspace(B, T, prm.n, prm.p, 2, obs.dtype, obs.device)
    hist = []

    def both(x1, x2):
        q = torch.as_tensor(np.exp([x1, x2]), dtype=obs.dtype).to(obs.device)
        nll, _ = dev.kf_nll(obs, prm, B, T, q_scale=q, workspace=ws)
        tot = allreduce_sum_(nll_sums(nll)).cpu().numpy()
        hist.extend([(float(np.exp(x1)), float(tot[0])), (float(np.exp(x2)), float(tot[1]))])
        return float(tot[0]), float(tot[1])

    x1, x2 = b - phi * (b - a), a + phi * (b - a)
    f1, f2 = both(x1, x2)
    for _ in range(iters):
        if f1 <= f2:
            b, x2, f2 = x2, x1, f1
            x1 = b - phi * (b - a)
        else:
            a, x1, f1 = x1, x2, f2
            x2 = a + phi * (b - a)
        # (one of the two is a re-evaluation: it keeps the launch shape -- and the workspace -- fixed)
        f1, f2 = both(x1, x2)
    best = min(hist, key=lambda e: e[1])
    return best[0], hist


def zscore(obs, smth, prm, B, T, add_ev=True):
    """p-row records of z = (y - C ms)/sqrt(diag(C Vs C^T) [+ ev])."""
    z = dev.alloc_records(B, T, prm.p, obs.dtype, obs.device)
    fn = getattr(lib(), f"eksb_zscore_{dev._sfx(obs.dtype)}")
    check(fn(obs.data_ptr(), smth.data_ptr(), prm.C.data_ptr(), prm.shared_mask, B, T, prm.n, prm.p, int(bool(add_ev)),
             z.data_ptr(), torch.cuda.current_stream(obs.device).cuda_stream), "eksb_zscore")
    return z
