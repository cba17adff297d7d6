"""Device-resident (torch.Tensor) API over the libeksb C ABI.

PyTorch is plumbing here: it owns device memory and streams; every computation is a
hand-written sm_100a kernel in eks_b200/csrc reached through ctypes.  Tensors in the
"track-tile" layout (include/eksb.h) are flat 1-D tensors of n_tiles*T*rows*32 scalars.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from . import _lib
from ._lib import LANES, check, lib


def _sfx(dtype):
    if dtype == torch.float64:
        return "f64"
    if dtype == torch.float32:
        return "f32"
    raise TypeError(f"eks_b200 computes in float32 or float64, got {dtype}")


def _ptr(t):
    return None if t is None else t.data_ptr()


def _stream(device):
    return torch.cuda.current_stream(device).cuda_stream


def _need_cuda(t, name):
    if not t.is_cuda:
        raise RuntimeError(f"eks_b200: '{name}' must be a CUDA tensor (there is no CPU path)")
    if not t.is_contiguous():
        raise RuntimeError(f"eks_b200: '{name}' must be contiguous")


def num_tiles(B):
    return (B + LANES - 1) // LANES


def sigma(n):
    return n + n * (n + 1) // 2


def alloc_records(B, T, rows, dtype, device):
    return torch.empty(num_tiles(B) * T * rows * LANES, dtype=dtype, device=device)


@dataclass
class TrackParams:
    """Per-track (leading dim B) or shared (no leading dim) model matrices on the device.

    Names follow eks/ensemble_kalman.py:59: m0 (n,), S0 (n,n), C (p,n), A (n,n), Q (n,n).
    """
    m0: torch.Tensor
    S0: torch.Tensor
    C: torch.Tensor
    A: torch.Tensor
    Q: torch.Tensor
    a_mode: int = -1

    def __post_init__(self):
        self.p, self.n = int(self.C.shape[-2]), int(self.C.shape[-1])
        mask = 0
        for bit, t, nd in ((_lib.SHARED_M0, self.m0, 1), (_lib.SHARED_S0, self.S0, 2), (_lib.SHARED_C, self.C, 2),
                           (_lib.SHARED_A, self.A, 2), (_lib.SHARED_Q, self.Q, 2)):
            _need_cuda(t, "params")
            if t.dim() == nd:
                mask |= bit
            elif t.dim() != nd + 1:
                raise ValueError("parameter arrays must be (..) or (B, ..)")
        self.shared_mask = mask
        if self.a_mode < 0:
            A = self.A
            off = A - torch.diag_embed(torch.diagonal(A, dim1=-2, dim2=-1))
            self.a_mode = _lib.A_DIAG if not bool((off != 0).any()) else _lib.A_DENSE

    @property
    def fast_path(self):
        """True if (n,p) has a register-resident template instantiation."""
        return bool(lib().eksb_supported_dims(self.n, self.p))

    @property
    def dtype(self):
        return self.C.dtype

    @classmethod
    def from_numpy(cls, m0, S0, C, A, Q, dtype=torch.float64, device="cuda"):
        f = lambda a: torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(device).contiguous()
        A = np.asarray(A)
        off = A - A * np.eye(A.shape[-1])
        return cls(f(m0), f(S0), f(C), f(A), f(Q), a_mode=_lib.A_DIAG if not np.any(off != 0) else _lib.A_DENSE)


# ------------------------------------------------------------------ layout
def pack_rows(src, dst, rows, row0, pad=0.0):
    """src (B,T,R) -> rows [row0,row0+R) of the track-tile buffer dst."""
    _need_cuda(src, "src")
    B, T, R = src.shape
    fn = getattr(lib(), f"eksb_pack_rows_{_sfx(src.dtype)}")
    check(fn(_ptr(src), B, T, R, _ptr(dst), rows, row0, pad, _stream(src.device)), "eksb_pack_rows")
    return dst


def unpack_rows(src, rows, row0, B, T, R):
    """rows [row0,row0+R) of a track-tile buffer -> (B,T,R)."""
    out = torch.empty((B, T, R), dtype=src.dtype, device=src.device)
    fn = getattr(lib(), f"eksb_unpack_rows_{_sfx(src.dtype)}")
    check(fn(_ptr(src), rows, row0, B, T, R, _ptr(out), _stream(src.device)), "eksb_unpack_rows")
    return out


def pack_obs(y, ev):
    """y, ev (B,T,p) -> obs records (padding lanes: y=0, ev=1)."""
    B, T, p = y.shape
    obs = alloc_records(B, T, 2 * p, y.dtype, y.device)
    pack_rows(y, obs, 2 * p, 0, 0.0)
    pack_rows(ev, obs, 2 * p, p, 1.0)
    return obs


def expand_state(state, B, T, prm: TrackParams, want_pred=False):
    """state records -> mean (B,T,n), cov (B,T,n,n) [, pred (B,T,n,n) = the reference's shifted S]."""
    n = prm.n
    mean = torch.empty((B, T, n), dtype=state.dtype, device=state.device)
    cov = torch.empty((B, T, n, n), dtype=state.dtype, device=state.device)
    pred = torch.empty((B, T, n, n), dtype=state.dtype, device=state.device) if want_pred else None
    fn = getattr(lib(), f"eksb_expand_state_{_sfx(state.dtype)}")
    check(fn(_ptr(state), B, T, n, _ptr(prm.A), _ptr(prm.Q), _ptr(prm.S0), prm.shared_mask, _ptr(mean), _ptr(cov),
             _ptr(pred), _stream(state.device)), "eksb_expand_state")
    return (mean, cov, pred) if want_pred else (mean, cov)


def pack_state(mean, cov):
    """dense mean (B,T,n), cov (B,T,n,n) -> state records."""
    _need_cuda(mean, "mean"); _need_cuda(cov, "cov")
    B, T, n = mean.shape
    state = alloc_records(B, T, sigma(n), mean.dtype, mean.device)
    fn = getattr(lib(), f"eksb_pack_state_{_sfx(mean.dtype)}")
    check(fn(_ptr(mean), _ptr(cov), B, T, n, _ptr(state), _stream(mean.device)), "eksb_pack_state")
    return state


def project(state, prm: "TrackParams", B, T, rows=None):
    """ym records = C * (mean rows of a state record buffer)."""
    rows = sigma(prm.n) if rows is None else rows
    ym = alloc_records(B, T, prm.p, state.dtype, state.device)
    fn = getattr(lib(), f"eksb_project_{_sfx(state.dtype)}")
    check(fn(_ptr(state), rows, _ptr(prm.C), prm.shared_mask, B, T, prm.n, prm.p, _ptr(ym), _stream(state.device)),
          "eksb_project")
    return ym


# ------------------------------------------------------------------ A1 ensemble
def _mode(mode):
    if mode == "median":
        return _lib.MODE_MEDIAN
    if mode == "mean":
        return _lib.MODE_MEAN
    raise ValueError(f"{mode} averaging not supported")      # eks/ensemble_kalman.py:39


def ensemble(preds, mode="median"):
    """preds (M,T,K) -> (avg (T,K), var (T,K)); eks/ensemble_kalman.py:45-46."""
    md = _mode(mode)
    _need_cuda(preds, "preds")
    M, T, K = preds.shape
    avg = torch.empty((T, K), dtype=preds.dtype, device=preds.device)
    var = torch.empty_like(avg)
    fn = getattr(lib(), f"eksb_ensemble_{_sfx(preds.dtype)}")
    check(fn(_ptr(preds), M, T, K, md, _ptr(avg), _ptr(var), _stream(preds.device)), "eksb_ensemble")
    return avg, var


def ensemble_tiled(preds, B, p, mode="median", offset=None, out=None, max_var=None):
    """preds (M,T,B*p) -> obs records, y = avg - offset (offset (B,p) or None).
    max_var: optional [tile][T][32] buffer (setup_max_var_view) that receives max_j ev[t,j] for pca_setup(mx_ready=True)."""
    md = _mode(mode)
    _need_cuda(preds, "preds")
    M, T, K = preds.shape
    if K != B * p:
        raise ValueError("preds last dim must be B*p")
    obs = out if out is not None else alloc_records(B, T, 2 * p, preds.dtype, preds.device)
    if max_var is not None:
        fn = getattr(lib(), f"eksb_ensemble_tiled_mx_{_sfx(preds.dtype)}")
        check(fn(_ptr(preds), M, T, B, p, md, _ptr(offset), _ptr(obs), _ptr(max_var), _stream(preds.device)),
              "eksb_ensemble_tiled_mx")
        return obs
    fn = getattr(lib(), f"eksb_ensemble_tiled_{_sfx(preds.dtype)}")
    check(fn(_ptr(preds), M, T, B, p, md, _ptr(offset), _ptr(obs), _stream(preds.device)), "eksb_ensemble_tiled")
    return obs


def setup_workspace(B, T, p, dtype, device):
    """Workspace of pca_setup, allocated up front so that ensemble_tiled can write max_vars into it."""
    es = torch.empty(0, dtype=dtype).element_size()
    nbytes = lib().eksb_pca_setup_workspace_bytes(B, T, p, es)
    return torch.empty(max(1, nbytes // 8), dtype=torch.float64, device=device)


def setup_max_var_view(ws, B, T, p, dtype):
    """The max_vars region of a pca_setup workspace as a flat [tile][T][32] tensor of the obs dtype."""
    es = torch.empty(0, dtype=dtype).element_size()
    off = lib().eksb_pca_setup_mx_offset_bytes(B, T, p, es)
    n = num_tiles(B) * T * 32
    return ws.view(torch.uint8)[off:off + n * es].view(dtype)


# ------------------------------------------------------------------ view alignment
def interp_rows(ts_src, src, ts_q, flip=None):
    """src (S,Ts,K) sampled at ts_src (Ts) -> (S,Tq,K) at ts_q (Tq); flip (K): out = flip - value where != 0
    (multiview_pca_smoother.py:70-101, scipy interp1d arithmetic, bit-identical)."""
    _need_cuda(src, "src")
    S, Ts, K = src.shape
    Tq = ts_q.shape[0]
    out = torch.empty((S, Tq, K), dtype=src.dtype, device=src.device)
    fn = getattr(lib(), f"eksb_interp_rows_{_sfx(src.dtype)}")
    check(fn(_ptr(ts_src), _ptr(src), S, Ts, K, _ptr(ts_q), Tq, _ptr(flip), _ptr(out), _stream(src.device)),
          "eksb_interp_rows")
    return out


# ------------------------------------------------------------------ K8 parameter setup on the device
def pca_setup(obs, B, T, p, quantile, smooth_param, n=3, center=True, status=None, workspace=None, mx_ready=False):
    """Per-track (means, C, S0, Q) of the multi-camera smoother from UNcentred obs records
    (multiview_pca_smoother.py:684-731); with center=True the y rows of `obs` are centred in place.
    Returns (TrackParams with m0 = 0, A = I, means (B,p) tensor, status (B,) int32)."""
    _need_cuda(obs, "obs")
    dev, dt = obs.device, obs.dtype
    means = torch.empty((B, p), dtype=dt, device=dev)
    C = torch.empty((B, p, n), dtype=dt, device=dev)
    S0 = torch.empty((B, n, n), dtype=dt, device=dev)
    Q = torch.empty((B, n, n), dtype=dt, device=dev)
    nbytes = lib().eksb_pca_setup_workspace_bytes(B, T, p, obs.element_size())
    if mx_ready and workspace is None:
        raise ValueError("mx_ready needs the workspace ensemble_tiled wrote max_vars into")
    ws = workspace if workspace is not None else torch.empty(max(1, nbytes // 8), dtype=torch.float64, device=dev)
    if ws.numel() * 8 < nbytes:
        raise ValueError("pca_setup workspace too small")
    if status is None:
        status = torch.zeros(B, dtype=torch.int32, device=dev)
    fn = getattr(lib(), f"eksb_pca_setup_mx_{_sfx(dt)}")
    check(fn(_ptr(obs), B, T, n, p, float(quantile), float(smooth_param), int(bool(center)), int(bool(mx_ready)),
             _ptr(means), _ptr(C), _ptr(S0), _ptr(Q), _ptr(ws), nbytes, _ptr(status), _stream(dev)), "eksb_pca_setup")
    prm = TrackParams(torch.zeros(n, dtype=dt, device=dev), S0, C, torch.eye(n, dtype=dt, device=dev), Q,
                      a_mode=_lib.A_DIAG)
    return prm, means, status


# ------------------------------------------------------------------ routing workspace
def route_workspace(B, T, n, p, dtype, device):
    """Workspace for kf_forward(workspace=...) / rts_backward(workspace=...): with it the library sends batches
    that have too few track tiles for the sequential kernels through the parallel-in-time scan
    (eksb_route_workspace_bytes; None when this shape always runs the sequential kernels)."""
    es = torch.empty(0, dtype=dtype).element_size()
    nbytes = lib().eksb_route_workspace_bytes(B, T, n, p, es)
    if nbytes == 0:
        return None
    return torch.empty((nbytes + es - 1) // es, dtype=dtype, device=device)


def routes_to_scan(B, T, n, p, dtype):
    return lib().eksb_route_workspace_bytes(B, T, n, p, torch.empty(0, dtype=dtype).element_size()) > 0


# ------------------------------------------------------------------ A2 / A6 forward
def kf_forward(obs, prm: TrackParams, B, T, first=_lib.FIRST_PRIOR_UPDATE, q_scale=None, want_filt=True,
               want_nll=False, filt=None, status=None, y_offset=None, workspace=None):
    """Forward filter.  Returns (filt or None, nll (n_scale,B) float64 or None, status (B,) int32).
    y_offset: (B,p) per-track offsets subtracted from y as the frames are read (the camera means of
    multiview_pca_smoother.py:698-705 without a centring pass over the records); not with q_scale.
    workspace: route_workspace(...) tensor; lets the library pick the parallel-in-time kernels for few-tile batches.
    status bits: 1 = non-positive pivot, 2 = a non-positive ensemble variance (see include/eksb.h)."""
    _need_cuda(obs, "obs")
    dev, dt = obs.device, obs.dtype
    if prm.dtype != dt:
        raise TypeError("obs and params dtypes differ")
    n_scale = 0 if q_scale is None else int(q_scale.numel())
    if want_filt and filt is None:
        filt = alloc_records(B, T, sigma(prm.n), dt, dev)
    if not want_filt:
        filt = None
    nll = torch.empty((max(n_scale, 1), B), dtype=torch.float64, device=dev) if want_nll else None
    if status is None:
        status = torch.zeros(B, dtype=torch.int32, device=dev)
    if y_offset is not None:
        if q_scale is not None:
            raise ValueError("y_offset and q_scale cannot be combined")
        if y_offset.dtype != dt or tuple(y_offset.shape) != (B, prm.p) or not y_offset.is_contiguous():
            raise ValueError("y_offset must be a contiguous (B,p) tensor of the obs dtype")
    if workspace is not None and q_scale is None:
        if workspace.dtype != dt:
            raise TypeError("workspace dtype differs from obs")
        fn = getattr(lib(), f"eksb_kf_forward_ws_{_sfx(dt)}")
        check(fn(_ptr(obs), _ptr(y_offset), _ptr(prm.m0), _ptr(prm.S0), _ptr(prm.C), _ptr(prm.A), _ptr(prm.Q),
                 prm.shared_mask, B, T, prm.n, prm.p, prm.a_mode, first, _ptr(filt), _ptr(nll), _ptr(status),
                 _ptr(workspace), workspace.numel() * workspace.element_size(), _stream(dev)), "eksb_kf_forward_ws")
        return filt, nll, status
    if y_offset is not None:
        fn = getattr(lib(), f"eksb_kf_forward_offset_{_sfx(dt)}")
        check(fn(_ptr(obs), _ptr(y_offset), _ptr(prm.m0), _ptr(prm.S0), _ptr(prm.C), _ptr(prm.A), _ptr(prm.Q),
                 prm.shared_mask, B, T, prm.n, prm.p, prm.a_mode, first, _ptr(filt), _ptr(nll), _ptr(status),
                 _stream(dev)), "eksb_kf_forward_offset")
        return filt, nll, status
    fn = getattr(lib(), f"eksb_kf_forward_{_sfx(dt)}")
    check(fn(_ptr(obs), _ptr(prm.m0), _ptr(prm.S0), _ptr(prm.C), _ptr(prm.A), _ptr(prm.Q), prm.shared_mask, B, T,
             prm.n, prm.p, prm.a_mode, first, _ptr(q_scale), n_scale, _ptr(filt), _ptr(nll), _ptr(status),
             _stream(dev)), "eksb_kf_forward")
    return filt, nll, status


def nll_workspace(B, T, n, p, n_scale, dtype, device):
    """Workspace tensor for kf_nll (None when the plain launch is the route for this shape)."""
    nbytes = lib().eksb_nll_workspace_bytes(B, T, n, p, int(n_scale), torch.empty(0, dtype=dtype).element_size())
    return torch.empty(nbytes, dtype=torch.uint8, device=device) if nbytes else None


def kf_nll(obs, prm: TrackParams, B, T, q_scale=None, first=_lib.FIRST_PRIOR_UPDATE, y_offset=None, status=None,
           workspace="auto"):
    """A8: innovation NLL of the forward filter per track and per candidate q_scale[g] * Q -> (nll (max(G,1), B)
    float64, status (B,) int32).  No state records.  workspace: "auto" allocates what eksb_nll_workspace_bytes asks
    for (few tiles + long tracks run time-split, include/eksb.h), a nll_workspace(...) tensor reuses one across calls
    (an optimiser's loop), None forces the plain launch."""
    _need_cuda(obs, "obs")
    dev, dt = obs.device, obs.dtype
    if prm.dtype != dt:
        raise TypeError("obs and params dtypes differ")
    n_scale = 0 if q_scale is None else int(q_scale.numel())
    if q_scale is not None and (q_scale.dtype != dt or not q_scale.is_contiguous()):
        raise ValueError("q_scale must be a contiguous tensor of the obs dtype")
    if y_offset is not None and (y_offset.dtype != dt or tuple(y_offset.shape) != (B, prm.p) or not y_offset.is_contiguous()):
        raise ValueError("y_offset must be a contiguous (B,p) tensor of the obs dtype")
    nll = torch.empty((max(n_scale, 1), B), dtype=torch.float64, device=dev)
    if status is None:
        status = torch.zeros(B, dtype=torch.int32, device=dev)
    if isinstance(workspace, str):
        workspace = nll_workspace(B, T, prm.n, prm.p, n_scale, dt, dev)
    fn = getattr(lib(), f"eksb_kf_nll_ws_{_sfx(dt)}")
    check(fn(_ptr(obs), _ptr(y_offset), _ptr(prm.m0), _ptr(prm.S0), _ptr(prm.C), _ptr(prm.A), _ptr(prm.Q),
             prm.shared_mask, B, T, prm.n, prm.p, prm.a_mode, first, _ptr(q_scale), n_scale, _ptr(nll), _ptr(status),
             _ptr(workspace), 0 if workspace is None else workspace.numel() * workspace.element_size(), _stream(dev)),
          "eksb_kf_nll_ws")
    return nll, status


def kf_forward_generic(obs, prm: TrackParams, B, T, R=None, first=_lib.FIRST_PRIOR_UPDATE, status=None):
    """Literal-arithmetic forward filter (dense R (p,p)/(B,p,p), any n<=8, p<=16).  Returns (filt, status)."""
    _need_cuda(obs, "obs")
    dev, dt = obs.device, obs.dtype
    filt = alloc_records(B, T, sigma(prm.n), dt, dev)
    if status is None:
        status = torch.zeros(B, dtype=torch.int32, device=dev)
    r_shared = 1 if (R is None or R.dim() == 2) else 0
    fn = getattr(lib(), f"eksb_kf_forward_generic_{_sfx(dt)}")
    check(fn(_ptr(obs), _ptr(prm.m0), _ptr(prm.S0), _ptr(prm.C), _ptr(prm.A), _ptr(prm.Q), _ptr(R), r_shared,
             prm.shared_mask, B, T, prm.n, prm.p, first, _ptr(filt), _ptr(status), _stream(dev)),
          "eksb_kf_forward_generic")
    return filt, status


# ------------------------------------------------------------------ A4 / A5 backward
def rts_backward(filt, prm: TrackParams, B, T, inplace=False, want_ym=True, want_yv=False, want_cv=False,
                 smth=None, ym=None, yv=None, status=None, workspace=None):
    """RTS pass.  Returns dict(smth, ym, yv, cv, status); smth is filt itself when inplace.
    workspace: route_workspace(...) tensor (may be the one the forward call used)."""
    _need_cuda(filt, "filt")
    dev, dt = filt.device, filt.dtype
    n, p = prm.n, prm.p
    if smth is None:
        smth = filt if inplace else torch.empty_like(filt)
    if want_ym and ym is None:
        ym = alloc_records(B, T, p, dt, dev)
    if not want_ym:
        ym = None
    if want_yv and yv is None:
        yv = alloc_records(B, T, p, dt, dev)
    if not want_yv:
        yv = None
    cv = alloc_records(B, T, n * n, dt, dev).zero_() if want_cv else None
    if status is None:
        status = torch.zeros(B, dtype=torch.int32, device=dev)
    if workspace is not None:
        if workspace.dtype != dt:
            raise TypeError("workspace dtype differs from filt")
        fn = getattr(lib(), f"eksb_rts_backward_ws_{_sfx(dt)}")
        check(fn(_ptr(filt), _ptr(prm.C), _ptr(prm.A), _ptr(prm.Q), prm.shared_mask, B, T, n, p, prm.a_mode, _ptr(smth),
                 _ptr(ym), _ptr(yv), _ptr(cv), _ptr(status), _ptr(workspace),
                 workspace.numel() * workspace.element_size(), _stream(dev)), "eksb_rts_backward_ws")
        return dict(smth=smth, ym=ym, yv=yv, cv=cv, status=status)
    fn = getattr(lib(), f"eksb_rts_backward_{_sfx(dt)}")
    check(fn(_ptr(filt), _ptr(prm.C), _ptr(prm.A), _ptr(prm.Q), prm.shared_mask, B, T, n, p, prm.a_mode, _ptr(smth),
             _ptr(ym), _ptr(yv), _ptr(cv), _ptr(status), _stream(dev)), "eksb_rts_backward")
    return dict(smth=smth, ym=ym, yv=yv, cv=cv, status=status)


def rts_backward_generic(filt, prm: TrackParams, B, T, pred=None, want_ym=True, want_yv=False, want_cv=False,
                         status=None):
    """Literal-arithmetic RTS pass (LU solve per frame, any n<=8, p<=16, dense A).  pred: optional packed
    NS-row records with the caller's own S[t] (eks/ensemble_kalman.py:158 reads its S argument).
    Returns dict(smth, ym, yv, cv, status)."""
    _need_cuda(filt, "filt")
    dev, dt = filt.device, filt.dtype
    n, p = prm.n, prm.p
    smth = torch.empty_like(filt)
    ym = alloc_records(B, T, p, dt, dev) if want_ym else None
    yv = alloc_records(B, T, p, dt, dev) if want_yv else None
    cv = alloc_records(B, T, n * n, dt, dev).zero_() if want_cv else None
    if status is None:
        status = torch.zeros(B, dtype=torch.int32, device=dev)
    fn = getattr(lib(), f"eksb_rts_backward_generic_{_sfx(dt)}")
    check(fn(_ptr(filt), _ptr(pred), _ptr(prm.C), _ptr(prm.A), _ptr(prm.Q), prm.shared_mask, B, T, n, p, _ptr(smth),
             _ptr(ym), _ptr(yv), _ptr(cv), _ptr(status), _stream(dev)), "eksb_rts_backward_generic")
    return dict(smth=smth, ym=ym, yv=yv, cv=cv, status=status)


def pack_sym(dense):
    """dense (B,T,n,n) -> packed NS-row records (upper triangle of the symmetrised matrix, row-major)."""
    B, T, n, _ = dense.shape
    iu = torch.triu_indices(n, n, device=dense.device)
    sym = 0.5 * (dense + dense.transpose(-1, -2))
    rows = sym[:, :, iu[0], iu[1]].contiguous()
    ns = rows.shape[-1]
    rec = alloc_records(B, T, ns, dense.dtype, dense.device)
    return pack_rows(rows, rec, ns, 0, 0.0)


# ------------------------------------------------------------------ A7 block-tridiagonal solve
def btd_solve(obs, prm: TrackParams, B, T, want_ym=False, inplace=True, status=None, scan=None):
    """Block-tridiagonal Newton step.  prm.C is the observation matrix B, prm.Q the process noise E.
    Returns dict(sol [x | diag blocks of H^-1, state-record layout], ym, status).

    scan: None = parallel-in-time for a handful of long tracks (fewer track tiles than SMs, T >= 4096:
    config 5), False = block-Thomas elimination (eksb_btd_solve), True = force the scan.  The scan route uses
    the identity the solver is pinned by (x == RTS mean, diagonal blocks of H^-1 == Vs):
    the same system is solved by the chunked filter + smoother elements of eksb_kf_rts_scan."""
    _need_cuda(obs, "obs")
    dev, dt = obs.device, obs.dtype
    if scan is None:
        scan = routes_to_scan(B, T, prm.n, prm.p, dt)
    if scan:
        out = kf_rts_scan(obs, prm, B, T, smooth=True, inplace=True, want_ym=want_ym, status=status)
        return dict(sol=out["smth"], ym=out["ym"], status=out["status"])
    work = alloc_records(B, T, sigma(prm.n), dt, dev)
    sol = work if inplace else torch.empty_like(work)
    ym = alloc_records(B, T, prm.p, dt, dev) if want_ym else None
    if status is None:
        status = torch.zeros(B, dtype=torch.int32, device=dev)
    fn = getattr(lib(), f"eksb_btd_solve_{_sfx(dt)}")
    check(fn(_ptr(obs), _ptr(prm.m0), _ptr(prm.S0), _ptr(prm.A), _ptr(prm.C), _ptr(prm.Q), prm.shared_mask, B, T,
             prm.n, prm.p, prm.a_mode, _ptr(work), _ptr(sol), _ptr(ym), _ptr(status), _stream(dev)), "eksb_btd_solve")
    return dict(sol=sol, ym=ym, status=status)


# ------------------------------------------------------------------ K4 parallel-in-time scan
def kf_rts_scan(obs, prm: TrackParams, B, T, chunk=0, smooth=True, inplace=True, want_ym=True, status=None,
                filt=None, ym=None, workspace=None):
    """Filter (+ RTS smoother) by the exact chunked scan over time; for few, very long tracks.
    Returns dict(filt, smth, ym, status, chunk).  filt / ym / workspace may be caller-provided."""
    _need_cuda(obs, "obs")
    dev, dt = obs.device, obs.dtype
    es = obs.element_size()
    nbytes = lib().eksb_scan_workspace_bytes(B, T, prm.n, chunk, es)
    if workspace is not None:
        if workspace.numel() * es < nbytes:
            raise ValueError("scan workspace too small")
        ws = workspace
    else:
        ws = torch.empty(max(1, nbytes // es), dtype=dt, device=dev)
    if filt is None:
        filt = alloc_records(B, T, sigma(prm.n), dt, dev)
    smth = (filt if inplace else torch.empty_like(filt)) if smooth else None
    if smooth and want_ym and ym is None:
        ym = alloc_records(B, T, prm.p, dt, dev)
    if not (smooth and want_ym):
        ym = None
    if status is None:
        status = torch.zeros(B, dtype=torch.int32, device=dev)
    fn = getattr(lib(), f"eksb_kf_rts_scan_{_sfx(dt)}")
    check(fn(_ptr(obs), _ptr(prm.m0), _ptr(prm.S0), _ptr(prm.C), _ptr(prm.A), _ptr(prm.Q), prm.shared_mask, B, T, prm.n,
             prm.p, prm.a_mode, chunk, _ptr(ws), nbytes, _ptr(filt), _ptr(smth), _ptr(ym), _ptr(status), _stream(dev)),
          "eksb_kf_rts_scan")
    return dict(filt=filt, smth=smth, ym=ym, status=status, chunk=int(lib().eksb_scan_chunk_len(B, T, prm.n, chunk, es)))
