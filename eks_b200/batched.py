"""Batched host-facing entry point: smooth many tracks in one call.

The reference smooths one track per ``filtering_pass`` + ``smooth_backward`` call pair inside
Python loops (scripts/multicam_example.py:106, test_script.py:43-63).  ``smooth_tracks`` is the
same computation for B tracks at once with HOST arrays in and out: track blocks are streamed
host -> device -> host through three CUDA streams per GPU (H2D copy, kernels, D2H copy) so PCIe
transfers overlap the kernels, and ``devices=[...]`` spreads the blocks of ONE host batch over
several GPUs from a single process (tracks are independent: no collective).  Everything numeric
runs in libeksb (no CPU fallback).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib, device as dev

#: what each output name holds; the reference smoothers only consume ``ym`` = C ms
#: (eks/multiview_pca_smoother.py:745,761-762) and compute-but-drop ``C Vs C^T`` (:746), whose diagonal is ``yv``.
OUTPUTS = ("ms", "Vs", "ym", "yv")


def _as_host_tensor(a, dtype):
    t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
    if t.dtype != dtype:
        t = t.to(dtype)
    return t.contiguous()


def alloc_outputs(B, T, n, p, dtype=torch.float64, pin=True, outputs=("ms", "Vs", "ym")):
    """Pre-allocated (optionally pinned) host outputs for ``smooth_tracks(out=...)``."""
    mk = lambda *s: torch.empty(s, dtype=dtype, pin_memory=pin)
    shapes = dict(ms=(B, T, n), Vs=(B, T, n * (n + 1) // 2), ym=(B, T, p), yv=(B, T, p))
    out = {k: mk(*shapes[k]) for k in outputs}
    out["status"] = torch.zeros(B, dtype=torch.int32, pin_memory=pin)
    return out


class _DeviceLane:
    """Streams, ring of device buffers and uploaded parameters of one GPU."""

    def __init__(self, d, blk, T, n, p, dtype, outputs, par, base_nd, lo, hi, use_ws, force_ws):
        self.d, self.lo, self.hi = d, lo, hi
        sg = dev.sigma(n)
        with torch.cuda.device(d):
            self.s_in, self.s_k, self.s_out = (torch.cuda.Stream(d) for _ in range(3))
            # parameters of this GPU's shard, uploaded once (a per-block upload from pageable memory would
            # synchronise the issuing thread once per block)
            self.par = [torch.as_tensor(np.ascontiguousarray(a if a.ndim == nd else a[lo:hi])).to(dtype).to(d)
                        for a, nd in zip(par, base_nd)]
            nblk = (hi - lo + blk - 1) // blk
            # scan workspace: not monotone in the track count (fewer tiles -> more chunks), so size it for the
            # full block AND the ragged last block of this shard
            last = (hi - lo) - (nblk - 1) * blk
            es = torch.empty(0, dtype=dtype).element_size()
            ws_bytes = max(_lib.lib().eksb_route_workspace_bytes(b, T, n, p, es) for b in {blk, last}) if use_ws else 0
            if force_ws:
                ws_bytes = max(ws_bytes, max(_lib.lib().eksb_scan_workspace_bytes(b, T, n, 0, es) for b in {blk, last}))
            self.bufs = []
            for _ in range(min(3, max(1, nblk))):
                bf = dict(
                    ws=torch.empty((ws_bytes + es - 1) // es, dtype=dtype, device=d) if ws_bytes else None,
                    y=torch.empty((blk, T, p), dtype=dtype, device=d), ev=torch.empty((blk, T, p), dtype=dtype, device=d),
                    obs=dev.alloc_records(blk, T, 2 * p, dtype, d), st=dev.alloc_records(blk, T, sg, dtype, d),
                    status=torch.empty(blk, dtype=torch.int32, device=d),
                    loaded=torch.cuda.Event(), computed=torch.cuda.Event(), drained=torch.cuda.Event())
                if "ym" in outputs:
                    bf["ymr"] = dev.alloc_records(blk, T, p, dtype, d)
                    bf["ym"] = torch.empty((blk, T, p), dtype=dtype, device=d)
                if "yv" in outputs:
                    bf["yvr"] = dev.alloc_records(blk, T, p, dtype, d)
                    bf["yv"] = torch.empty((blk, T, p), dtype=dtype, device=d)
                if "ms" in outputs:
                    bf["ms"] = torch.empty((blk, T, n), dtype=dtype, device=d)
                if "Vs" in outputs:
                    bf["Vs"] = torch.empty((blk, T, sg - n), dtype=dtype, device=d)
                self.bufs.append(bf)
            # the buffers were allocated on the caller's current stream; the side streams start after it
            ready = torch.cuda.Event()
            ready.record(torch.cuda.current_stream(d))
            for s in (self.s_in, self.s_k, self.s_out):
                s.wait_event(ready)
        self.next_block = 0

    def params(self, b0, b1, base_nd):
        sl = [t if t.dim() == nd else t[b0 - self.lo:b1 - self.lo] for t, nd in zip(self.par, base_nd)]
        return dev.TrackParams(*sl, a_mode=self.a_mode)


def smooth_tracks(y, ev, m0, S0, C, A, Q, *, dtype=torch.float64, device=None, devices=None, block_tracks=None,
                  out=None, outputs=("ym",), counters=None, scan=None):
    """Kalman filter + RTS smoother for B independent tracks (A2 + A4 + A5 of SURVEY section 8a).

    y, ev : host (B,T,p) arrays/tensors -- centred ensemble means and ensemble variances (pinned tensors make the
    copies asynchronous).  m0 (n,)|(B,n), S0 (n,n)|(B,n,n), C (p,n)|(B,p,n), A, Q (n,n)|(B,n,n): as in
    eks/ensemble_kalman.py:59 (a leading B dimension makes a parameter per-track).
    outputs: which results travel back to the host, any of
        "ym" (B,T,p) = C ms            -- all the reference smoothers consume (multiview_pca_smoother.py:745,761-762)
        "yv" (B,T,p) = diag(C Vs C^T)  -- the diagonal of :746
        "ms" (B,T,n), "Vs" (B,T,n(n+1)/2) [upper triangle, row-major] -- the smoothed latents themselves.
    Returns a dict of host tensors for those names plus status (B,) int32 [bit 1 = non-positive pivot, the
    reference's LinAlgError].  Tracks with a non-positive ensemble variance are re-run through the literal
    p x p LU kernel (the information-form fast path divides by every variance; the reference does not).
    devices: list of CUDA device indices; the blocks of the batch are dealt out over them (contiguous shard per GPU),
    all driven from this process.  block_tracks: tracks per streamed block; None = about 0.8 GB of input per block
    (128 tracks at 100 000 frames, p = 4, f64: measured best on B200, the exposed head and tail of the copy pipeline
    shrink with the block while the time-split kernels keep few-tile blocks on the roofline), at most 1024.  scan: None = the library's routing (parallel-in-time kernels when a block has
    too few tiles for the sequential ones), True / False = force.
    """
    if not torch.cuda.is_available():
        raise RuntimeError("eks_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    outputs = tuple(outputs)
    for k in outputs:
        if k not in OUTPUTS:
            raise ValueError(f"unknown output {k!r}; choose from {OUTPUTS}")
    if not outputs:
        raise ValueError("outputs is empty")
    if devices is None:
        d0 = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        dlist = [d0]
    else:
        dlist = [torch.device("cuda", int(i)) for i in devices]
        if not dlist:
            raise ValueError("devices is empty")
    y = _as_host_tensor(y, dtype)
    ev = _as_host_tensor(ev, dtype)
    B, T, p = y.shape
    par = [np.asarray(a, dtype=np.float64) for a in (m0, S0, C, A, Q)]
    n = par[2].shape[-1]
    base_nd = (1, 2, 2, 2, 2)
    if out is None:
        out = alloc_outputs(B, T, n, p, dtype, pin=False, outputs=outputs)
    for k in outputs:
        if k not in out:
            raise ValueError(f"out has no buffer for requested output {k!r}")
    sg = dev.sigma(n)
    fast = bool(_lib.lib().eksb_supported_dims(n, p))
    if not fast:
        raise _lib.EksbError(f"smooth_tracks: (n={n}, p={p}) has no batched kernel; supported {_lib.SUPPORTED_DIMS}")
    Aoff = par[3] - par[3] * np.eye(n)
    a_mode = _lib.A_DIAG if not np.any(Aoff != 0) else _lib.A_DENSE
    ndev = min(len(dlist), max(1, (B + 31) // 32))
    dlist = dlist[:ndev]
    per_dev = (B + ndev - 1) // ndev
    per_dev = (per_dev + 31) // 32 * 32                       # shard boundaries on tile boundaries
    if block_tracks is None:
        block_tracks = min(1024, max(32, int(round(0.8e9 / (T * 2 * p * y.element_size()) / 32)) * 32))
    blk = max(32, (min(block_tracks, per_dev) + 31) // 32 * 32)
    want_yv = "yv" in outputs
    # forcing: scan=True routes through the library's workspace API with EKSB_ROUTE honoured there; scan=False = no workspace
    use_ws = scan is not False and not want_yv
    lanes = []
    for i, d in enumerate(dlist):
        lo, hi = min(B, i * per_dev), min(B, (i + 1) * per_dev)
        if hi > lo:
            ln = _DeviceLane(d, blk, T, n, p, dtype, outputs, par, base_nd, lo, hi, use_ws, scan is True)
            ln.a_mode = a_mode
            lanes.append(ln)
    launches = 0
    force = None
    if scan is True:
        import os
        force = os.environ.get("EKSB_ROUTE")
        os.environ["EKSB_ROUTE"] = "scan"
    sfx = dev._sfx(dtype)
    unpack = getattr(_lib.lib(), f"eksb_unpack_rows_{sfx}")
    try:
        more = True
        while more:                                            # deal blocks round-robin so every GPU's queues stay fed
            more = False
            for ln in lanes:
                b0 = ln.lo + ln.next_block * blk
                if b0 >= ln.hi:
                    continue
                more = True
                b1 = min(ln.hi, b0 + blk)
                nb = b1 - b0
                bf = ln.bufs[ln.next_block % len(ln.bufs)]
                ln.next_block += 1
                with torch.cuda.device(ln.d):
                    with torch.cuda.stream(ln.s_in):
                        ln.s_in.wait_event(bf["computed"])       # previous occupant's kernels are done with y/ev
                        bf["y"][:nb].copy_(y[b0:b1], non_blocking=True)
                        bf["ev"][:nb].copy_(ev[b0:b1], non_blocking=True)
                        bf["loaded"].record(ln.s_in)
                    prm = ln.params(b0, b1, base_nd)
                    with torch.cuda.stream(ln.s_k):
                        ln.s_k.wait_event(bf["loaded"])
                        ln.s_k.wait_event(bf["drained"])         # previous occupant's outputs have left the device
                        bf["status"].zero_()
                        dev.pack_rows(bf["y"][:nb], bf["obs"], 2 * p, 0, 0.0)
                        dev.pack_rows(bf["ev"][:nb], bf["obs"], 2 * p, p, 1.0)
                        ws = bf["ws"]
                        dev.kf_forward(bf["obs"], prm, nb, T, filt=bf["st"], status=bf["status"][:nb], workspace=ws)
                        dev.rts_backward(bf["st"], prm, nb, T, inplace=True, want_ym="ym" in outputs, ym=bf.get("ymr"),
                                         want_yv=want_yv, yv=bf.get("yvr"), status=bf["status"][:nb], workspace=ws)
                        launches += 4 + (6 if (ws is not None) else 2)
                        for name, src, rows, r0, R in (("ms", "st", sg, 0, n), ("Vs", "st", sg, n, sg - n),
                                                       ("ym", "ymr", p, 0, p), ("yv", "yvr", p, 0, p)):
                            if name in outputs:
                                dev.check(unpack(bf[src].data_ptr(), rows, r0, nb, T, R, bf[name].data_ptr(),
                                                 ln.s_k.cuda_stream), "eksb_unpack_rows")
                                launches += 1
                        bf["computed"].record(ln.s_k)
                    with torch.cuda.stream(ln.s_out):
                        ln.s_out.wait_event(bf["computed"])
                        for name in outputs:
                            out[name][b0:b1].copy_(bf[name][:nb], non_blocking=True)
                        out["status"][b0:b1].copy_(bf["status"][:nb], non_blocking=True)
                        bf["drained"].record(ln.s_out)
    finally:
        if scan is True:
            import os
            if force is None:
                os.environ.pop("EKSB_ROUTE", None)
            else:
                os.environ["EKSB_ROUTE"] = force
    for ln in lanes:
        ln.s_out.synchronize()
        ln.s_k.synchronize()
    launches += _rerun_nonpositive_variance(y, ev, par, base_nd, out, outputs, dtype, lanes[0].d, a_mode)
    if counters is not None:
        counters["launches"] = counters.get("launches", 0) + launches
        counters["h2d_bytes"] = counters.get("h2d_bytes", 0) + 2 * y.numel() * y.element_size()
        counters["d2h_bytes"] = counters.get("d2h_bytes", 0) + sum(out[k].numel() * out[k].element_size()
                                                                   for k in outputs + ("status",))
        counters["devices"] = len(lanes)
    return out


def _rerun_nonpositive_variance(y, ev, par, base_nd, out, outputs, dtype, d, a_mode):
    """Tracks the fast kernels flagged with EKSB_STATUS_NONPOS_VAR (an ensemble variance <= 0 in some frame: every
    model agreed to the last bit) are filtered again with the literal p x p LU kernel, which has the reference's
    semantics there (eks/ensemble_kalman.py:110-117), then smoothed by the ordinary RTS kernel."""
    st = out["status"]
    bad = torch.nonzero((st & _lib.STATUS_NONPOS_VAR) != 0).flatten()
    if bad.numel() == 0:
        return 0
    idx = bad.tolist()
    nb = len(idx)
    T, p = y.shape[1], y.shape[2]
    n = par[2].shape[-1]
    sg = dev.sigma(n)
    with torch.cuda.device(d):
        yb = y[bad].to(d)
        evb = ev[bad].to(d)
        sl = [a if a.ndim == nd else a[idx] for a, nd in zip(par, base_nd)]
        prm = dev.TrackParams.from_numpy(*sl, dtype=dtype, device=d)
        obs = dev.pack_obs(yb, evb)
        filt, status = dev.kf_forward_generic(obs, prm, nb, T)
        res = dev.rts_backward(filt, prm, nb, T, inplace=True, want_ym="ym" in outputs, want_yv="yv" in outputs,
                               status=status)
        if "ms" in outputs:
            out["ms"][bad] = dev.unpack_rows(res["smth"], sg, 0, nb, T, n).cpu()
        if "Vs" in outputs:
            out["Vs"][bad] = dev.unpack_rows(res["smth"], sg, n, nb, T, sg - n).cpu()
        if "ym" in outputs:
            out["ym"][bad] = dev.unpack_rows(res["ym"], p, 0, nb, T, p).cpu()
        if "yv" in outputs:
            out["yv"][bad] = dev.unpack_rows(res["yv"], p, 0, nb, T, p).cpu()
        st[bad] = status.cpu() & _lib.STATUS_PIVOT
    return 6
