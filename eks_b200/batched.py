"""Batched host-facing entry point: smooth many tracks in one call.

The reference smooths one track per ``filtering_pass`` + ``smooth_backward`` call pair inside
Python loops (scripts/multicam_example.py:106, test_script.py:43-63).  ``smooth_tracks`` is the
same computation for B tracks at once with HOST arrays in and out: track blocks are streamed
host -> device -> host through three CUDA streams (H2D copy, kernels, D2H copy) so PCIe
transfers overlap the kernels.  Everything numeric runs in libeksb (no CPU fallback).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib, device as dev


def _as_host_tensor(a, dtype):
    t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
    if t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


def alloc_outputs(B, T, n, p, dtype=torch.float64, pin=True):
    """Pre-allocated (optionally pinned) host outputs for ``smooth_tracks(out=...)``."""
    mk = lambda *s: torch.empty(s, dtype=dtype, pin_memory=pin)
    return dict(ms=mk(B, T, n), Vs=mk(B, T, n * (n + 1) // 2), ym=mk(B, T, p),
                status=torch.zeros(B, dtype=torch.int32, pin_memory=pin))


def smooth_tracks(y, ev, m0, S0, C, A, Q, *, dtype=torch.float64, device=None, block_tracks=1024, out=None,
                  counters=None, scan=None):
    """Kalman filter + RTS smoother for B independent tracks (A2 + A4 + A5 of SURVEY section 8a).

    y, ev : host (B,T,p) arrays/tensors -- centred ensemble means and ensemble variances.
    m0 (n,)|(B,n), S0 (n,n)|(B,n,n), C (p,n)|(B,p,n), A, Q (n,n)|(B,n,n): as in
    eks/ensemble_kalman.py:59 (a leading B dimension makes a parameter per-track).
    Returns host tensors: ms (B,T,n), Vs (B,T,n(n+1)/2) [upper triangle, row-major],
    ym (B,T,p) = C ms, status (B,) [1 = non-positive pivot, the reference's LinAlgError].
    scan: None = choose per block size (parallel-in-time scan when a block has fewer tiles than SMs),
    True / False = force.
    """
    if not torch.cuda.is_available():
        raise RuntimeError("eks_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    d = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    y = _as_host_tensor(y, dtype)
    ev = _as_host_tensor(ev, dtype)
    B, T, p = y.shape
    par = [np.asarray(a, dtype=np.float64) for a in (m0, S0, C, A, Q)]
    n = par[2].shape[-1]
    base_nd = (1, 2, 2, 2, 2)
    if out is None:
        out = alloc_outputs(B, T, n, p, dtype, pin=False)
    sg = dev.sigma(n)
    blk = max(32, (min(block_tracks, B) + 31) // 32 * 32)
    nblk = (B + blk - 1) // blk
    s_in, s_k, s_out = (torch.cuda.Stream(d) for _ in range(3))
    nbuf = 3
    # sequential kernels need >= ~1 tile per SM to be worth their per-frame latency; below that use the scan
    use_scan = (scan if scan is not None else
                (dev.num_tiles(blk) < 148 and T >= 4096 and n <= 4 and bool(_lib.lib().eksb_supported_dims(n, p))))
    es = torch.empty(0, dtype=dtype).element_size()
    ws_elems = max(1, _lib.lib().eksb_scan_workspace_bytes(blk, T, n, 0, es) // es) if use_scan else 1
    bufs = []
    for _ in range(min(nbuf, nblk)):
        bufs.append(dict(
            ws=torch.empty(ws_elems, dtype=dtype, device=d),
            y=torch.empty((blk, T, p), dtype=dtype, device=d), ev=torch.empty((blk, T, p), dtype=dtype, device=d),
            obs=dev.alloc_records(blk, T, 2 * p, dtype, d), st=dev.alloc_records(blk, T, sg, dtype, d),
            ymr=dev.alloc_records(blk, T, p, dtype, d),
            ms=torch.empty((blk, T, n), dtype=dtype, device=d), Vs=torch.empty((blk, T, sg - n), dtype=dtype, device=d),
            ym=torch.empty((blk, T, p), dtype=dtype, device=d), status=torch.zeros(blk, dtype=torch.int32, device=d),
            loaded=torch.cuda.Event(), computed=torch.cuda.Event(), drained=torch.cuda.Event()))
    launches = 0
    keep = []          # parameter tensors stay referenced until all streams drain (cross-stream reuse)
    for i in range(nblk):
        b0, b1 = i * blk, min(B, (i + 1) * blk)
        nb = b1 - b0
        bf = bufs[i % len(bufs)]
        with torch.cuda.stream(s_in):
            s_in.wait_event(bf["computed"])          # previous occupant's kernels are done with y/ev
            bf["y"][:nb].copy_(y[b0:b1], non_blocking=True)
            bf["ev"][:nb].copy_(ev[b0:b1], non_blocking=True)
            bf["loaded"].record(s_in)
        sl = [a if a.ndim == nd else a[b0:b1] for a, nd in zip(par, base_nd)]
        prm = dev.TrackParams.from_numpy(*sl, dtype=dtype, device=d)
        keep.append(prm)
        with torch.cuda.stream(s_k):
            s_k.wait_event(bf["loaded"])
            s_k.wait_event(bf["drained"])            # previous occupant's outputs have left the device
            bf["status"].zero_()
            yb, evb = bf["y"][:nb], bf["ev"][:nb]
            dev.pack_rows(yb, bf["obs"], 2 * p, 0, 0.0)
            dev.pack_rows(evb, bf["obs"], 2 * p, p, 1.0)
            if use_scan:
                # few tiles, long tracks: the sequential kernels would run at one warp's pace, so
                # cut time into chunks (exact parallel-in-time scan, eksb_kf_rts_scan)
                dev.kf_rts_scan(bf["obs"], prm, nb, T, filt=bf["st"], ym=bf["ymr"], status=bf["status"][:nb],
                                workspace=bf["ws"])
                launches += 4                      # 6 scan kernels instead of 2
            else:
                dev.kf_forward(bf["obs"], prm, nb, T, filt=bf["st"], status=bf["status"][:nb])
                dev.rts_backward(bf["st"], prm, nb, T, inplace=True, ym=bf["ymr"], status=bf["status"][:nb])
            for name, rows, r0, R in (("ms", sg, 0, n), ("Vs", sg, n, sg - n)):
                fn = getattr(_lib.lib(), f"eksb_unpack_rows_{dev._sfx(dtype)}")
                dev.check(fn(bf["st"].data_ptr(), rows, r0, nb, T, R, bf[name].data_ptr(), s_k.cuda_stream), "eksb_unpack_rows")
            fn = getattr(_lib.lib(), f"eksb_unpack_rows_{dev._sfx(dtype)}")
            dev.check(fn(bf["ymr"].data_ptr(), p, 0, nb, T, p, bf["ym"].data_ptr(), s_k.cuda_stream), "eksb_unpack_rows")
            bf["computed"].record(s_k)
            launches += 7
        with torch.cuda.stream(s_out):
            s_out.wait_event(bf["computed"])
            out["ms"][b0:b1].copy_(bf["ms"][:nb], non_blocking=True)
            out["Vs"][b0:b1].copy_(bf["Vs"][:nb], non_blocking=True)
            out["ym"][b0:b1].copy_(bf["ym"][:nb], non_blocking=True)
            out["status"][b0:b1].copy_(bf["status"][:nb], non_blocking=True)
            bf["drained"].record(s_out)
    s_out.synchronize()
    s_k.synchronize()
    if counters is not None:
        counters["launches"] = counters.get("launches", 0) + launches
        counters["h2d_bytes"] = counters.get("h2d_bytes", 0) + 2 * y.numel() * y.element_size()
        counters["d2h_bytes"] = counters.get("d2h_bytes", 0) + sum(out[k].numel() * out[k].element_size()
                                                                   for k in ("ms", "Vs", "ym", "status"))
    return out
