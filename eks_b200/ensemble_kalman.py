"""Drop-in for ``eks/ensemble_kalman.py`` of erialc-cal/eks: same names, positional
signatures, dtypes, shapes and side effects -- computed by sm_100a CUDA kernels.

    from eks_b200.ensemble_kalman import ensemble, filtering_pass, kalman_dot, smooth_backward

numpy (host) in, numpy out; host<->device copies happen inside each call.  There is no CPU
fallback: without a CUDA device / libeksb.so these functions raise.  For many tracks use the
batched device API (eks_b200.device / eks_b200.batched) instead of looping over these.
"""
from __future__ import annotations

from collections import defaultdict

import numpy as np
import torch

from . import _lib, device as dev

_DT = torch.float64   # the reference is float64 throughout


def _cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("eks_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _to_dev(a, d):
    return torch.as_tensor(np.ascontiguousarray(a, dtype=np.float64)).to(d)


def _raise_if_singular(status):
    """A non-positive pivot anywhere in the batch is the reference's numpy.linalg.LinAlgError."""
    if bool(((status & _lib.STATUS_PIVOT) != 0).any().item()):
        raise np.linalg.LinAlgError("Singular matrix")


def _nonpositive_variance(status):
    """True if a fast kernel met an ensemble variance <= 0 (it cannot divide by it; the literal LU kernel can)."""
    return bool(((status & _lib.STATUS_NONPOS_VAR) != 0).any().item())


def ensemble(markers_list, keys, mode='median'):
    """eks/ensemble_kalman.py:4-57.  Same 6-tuple: ensemble_preds (T,K), ensemble_vars (T,K),
    ensemble_stacks (M,T,K) [memory order K,T,M like the reference], and the three dicts."""
    if mode not in ('median', 'mean'):
        raise ValueError(f"{mode} averaging not supported")
    d = _cuda()
    M, T, K = len(markers_list), markers_list[0].shape[0], len(keys)
    base = np.empty((K, T, M))                       # [K][T][M]  (reference :54-56)
    for m in range(M):
        for k, key in enumerate(keys):
            base[k, :, m] = np.asarray(markers_list[m][key], dtype=np.float64)
    stacks = base.T                                  # logical (M,T,K)
    preds_d = torch.as_tensor(np.ascontiguousarray(stacks)).to(d)   # [M][T][K] for coalesced reads
    avg, var = dev.ensemble(preds_d, mode)
    ensemble_preds = avg.cpu().numpy()
    ensemble_vars = var.cpu().numpy()
    keypoints_avg_dict, keypoints_var_dict = {}, {}
    keypoints_stack_dict = defaultdict(dict)
    for k, key in enumerate(keys):
        keypoints_avg_dict[key] = ensemble_preds[:, k].copy()
        keypoints_var_dict[key] = ensemble_vars[:, k].copy()
        for m in range(M):
            keypoints_stack_dict[m][key] = base[k, :, m]
    return ensemble_preds, ensemble_vars, stacks, keypoints_avg_dict, keypoints_var_dict, keypoints_stack_dict


def _params(d, m0, S0, C, A, Q):
    return dev.TrackParams.from_numpy(m0, S0, C, A, Q, dtype=_DT, device=d)


def filtering_pass(y, m0, S0, C, R, A, Q, ensemble_vars):
    """eks/ensemble_kalman.py:59-107.  Returns mf (T,n), Vf (T,n,n), S (T,n,n) with the
    reference's layout (S[t] = predicted covariance of frame t+1, S[T-1] = 0) and, like the
    reference, leaves diag(R) overwritten with ensemble_vars[T-1]."""
    d = _cuda()
    y = np.asarray(y, dtype=np.float64)
    ev = np.asarray(ensemble_vars, dtype=np.float64)
    T, p = y.shape
    n = np.asarray(m0).shape[0]
    obs = dev.pack_obs(_to_dev(y[None], d), _to_dev(ev[None], d))
    Rm = np.asarray(R, dtype=np.float64)
    dense_r = np.any((Rm - np.diag(np.diagonal(Rm))) != 0)
    status = torch.zeros(1, dtype=torch.int32, device=d)
    prm = _params(d, m0, S0, C, A, Q)
    if dense_r or not prm.fast_path:
        # literal p x p LU path: dense R off-diagonals (:112) or an (n,p) without a template kernel
        filt, status = dev.kf_forward_generic(obs, prm, 1, T, R=_to_dev(Rm, d), status=status)
    else:
        filt, _, status = dev.kf_forward(obs, prm, 1, T, status=status)
        if _nonpositive_variance(status):
            # every ensemble member agreed to the last bit in some frame: R has a zero on its diagonal, which the
            # reference's R + C S C^T solve (:112-113) takes in its stride; so does the literal kernel
            filt, status = dev.kf_forward_generic(obs, prm, 1, T, R=_to_dev(Rm, d))
    mean, cov, pred = dev.expand_state(filt, 1, T, prm, want_pred=True)
    mf, Vf, S = mean[0].cpu().numpy(), cov[0].cpu().numpy(), pred[0].cpu().numpy()
    _raise_if_singular(status)
    idx = np.arange(p)
    R[idx, idx] = ev[T - 1]                                  # side effect of :88-89,:99-100
    return mf, Vf, S


def kalman_dot(array, V, C, R):
    """eks/ensemble_kalman.py:110-117:  V C^T (R + C V C^T)^-1 array  (p x p LU solve on the device)."""
    d = _cuda()
    arr = np.asarray(array, dtype=np.float64)
    p, n = np.asarray(C).shape
    k = 1 if arr.ndim == 1 else arr.shape[1]
    out = torch.empty((n, k), dtype=_DT, device=d)
    status = torch.zeros(1, dtype=torch.int32, device=d)
    a_d, V_d, C_d, R_d = (_to_dev(x, d) for x in (arr.reshape(p, k), V, C, R))
    dev.check(_lib.lib().eksb_kalman_dot_f64(a_d.data_ptr(), k, V_d.data_ptr(), C_d.data_ptr(), R_d.data_ptr(), n, p,
                                             out.data_ptr(), status.data_ptr(),
                                             torch.cuda.current_stream(d).cuda_stream), "eksb_kalman_dot")
    _raise_if_singular(status)
    res = out.cpu().numpy()
    return res[:, 0] if arr.ndim == 1 else res


def smooth_backward(y, mf, Vf, S, A, Q, C):
    """eks/ensemble_kalman.py:120-164.  Returns ms (T,n), Vs (T,n,n), CV (T-1,n,n).
    Like the reference only y.shape[0] is used from y, and Q is not read when S is usable: the reference takes the
    predicted covariances from its S argument (:158-160).  When S is what filtering_pass returns (S[t] = A Vf[t] A^T
    + Q, :101) the register-resident kernel recomputes it on the device instead of uploading it; a caller-modified
    S, or an (n,p) without a template instantiation, goes through the literal LU kernel with S uploaded."""
    d = _cuda()
    T = np.asarray(y).shape[0]
    mf = np.asarray(mf, dtype=np.float64)
    Vf = np.asarray(Vf, dtype=np.float64)
    n = mf.shape[1]
    p = np.asarray(C).shape[0]
    prm = _params(d, np.zeros(n), np.eye(n), C, A, Q)
    filt = dev.pack_state(_to_dev(mf[None], d), _to_dev(Vf[None], d))
    S = None if S is None else np.asarray(S, dtype=np.float64)
    own_S = False
    if S is not None and T > 1:
        An, Qn = np.asarray(A, dtype=np.float64), np.asarray(Q, dtype=np.float64)
        want = An @ Vf[:T - 1] @ An.T + Qn
        own_S = not np.allclose(S[:T - 1], want, rtol=1e-9, atol=1e-12 * max(1.0, float(np.abs(want).max())))
    if own_S or not prm.fast_path:
        pred = dev.pack_sym(_to_dev(S[None], d)) if S is not None else None
        out = dev.rts_backward_generic(filt, prm, 1, T, pred=pred, want_ym=False, want_cv=True)
    else:
        out = dev.rts_backward(filt, prm, 1, T, want_ym=False, want_cv=True)
    _raise_if_singular(out["status"])
    mean, cov = dev.expand_state(out["smth"], 1, T, prm)
    cv = dev.unpack_rows(out["cv"], n * n, 0, 1, T, n * n)[0].cpu().numpy().reshape(T, n, n)
    return mean[0].cpu().numpy(), cov[0].cpu().numpy(), cv[:T - 1]
