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
from .ensemble_kalman import _cuda, _to_dev, _raise_if_singular, _params


def kalman_newton_recursive(y, mu0, S0, A, B, ensemble_vars, E, max_iter=1):
    """eks/newton_eks.py:115-148.  Returns q (T,r) for max_iter == 1, else (q, loss) with
    loss == 0 (the reference updates q in place through the alias at :129, so its loss is 0
    and every extra sweep recomputes the same q with a further-propagated P; only
    max_iter == 1 is used by the reference's callers and only it is supported here)."""
    if max_iter != 1:
        raise NotImplementedError("kalman_newton_recursive: only max_iter=1 (the value every reference caller uses)")
    d = _cuda()
    y = np.asarray(y, dtype=np.float64)
    ev = np.asarray(ensemble_vars, dtype=np.float64)
    T, p = y.shape
    r = np.asarray(mu0).shape[0]
    prm = _params(d, mu0, S0, B, A, E)
    obs = dev.pack_obs(_to_dev(y[None], d), _to_dev(ev[None], d))
    if prm.fast_path:
        filt, _, status = dev.kf_forward(obs, prm, 1, T, first=_lib.FIRST_PRIOR_ONLY)
    else:
        filt, status = dev.kf_forward_generic(obs, prm, 1, T, first=_lib.FIRST_PRIOR_ONLY)
    _raise_if_singular(status)
    return dev.unpack_rows(filt, dev.sigma(r), 0, 1, T, r)[0].cpu().numpy()
