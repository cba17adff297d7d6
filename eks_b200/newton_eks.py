"""Drop-in for the runnable part of ``eks/newton_eks.py``: ``kalman_newton_recursive``.

The Humphrey-West one-pass recursion (eks/newton_eks.py:134-143)
    P_t = inv(inv(E + A P A^T) + B^T D_t^-1 B),  q_t = A q_{t-1} - P_t B^T D_t^-1 (B A q_{t-1} - y_t)
is the information-form Kalman update the forward kernel already runs, so this is the
forward kernel with EKSB_FIRST_PRIOR_ONLY (q[0]=mu0, P0=inv(S0), y[0] unused; :128,:132).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib, device as dev
from .ensemble_kalman import _cuda, _to_dev, _raise_if_singular, _nonpositive_variance, _params


def kalman_newton_recursive(y, mu0, S0, A, B, ensemble_vars, E, max_iter=1):
    """eks/newton_eks.py:115-148.  Returns q (T,r) for max_iter == 1, else (q, loss).

    Every sweep of the reference restarts from q[0] = mu0 but carries P on from the end of the previous sweep
    (:132-143; `qnew = q` at :129 aliases the two arrays, so the update is in place and loss == 0 for every sweep).
    Here each sweep is one launch of the forward kernel with EKSB_FIRST_PRIOR_ONLY, whose prior precision is the
    inverse of the previous sweep's last filtered covariance."""
    max_iter = int(max_iter)
    if max_iter < 1:
        raise ValueError("max_iter must be >= 1")
    d = _cuda()
    y = np.asarray(y, dtype=np.float64)
    ev = np.asarray(ensemble_vars, dtype=np.float64)
    T, p = y.shape
    r = np.asarray(mu0).shape[0]
    obs = dev.pack_obs(_to_dev(y[None], d), _to_dev(ev[None], d))
    S0_it = np.asarray(S0, dtype=np.float64)
    for it in range(max_iter):
        prm = _params(d, mu0, S0_it, B, A, E)
        if prm.fast_path:
            filt, _, status = dev.kf_forward(obs, prm, 1, T, first=_lib.FIRST_PRIOR_ONLY)
            if _nonpositive_variance(status):
                filt, status = dev.kf_forward_generic(obs, prm, 1, T, first=_lib.FIRST_PRIOR_ONLY)
        else:
            filt, status = dev.kf_forward_generic(obs, prm, 1, T, first=_lib.FIRST_PRIOR_ONLY)
        _raise_if_singular(status)
        if it + 1 < max_iter:
            _, cov = dev.expand_state(filt, 1, T, prm)
            S0_it = np.linalg.inv(cov[0, T - 1].cpu().numpy())      # the kernel starts from P = inv(S0) (:132)
    q = dev.unpack_rows(filt, dev.sigma(r), 0, 1, T, r)[0].cpu().numpy()
    return q if max_iter == 1 else (q, np.zeros(max_iter))
