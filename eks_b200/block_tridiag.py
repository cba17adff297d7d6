"""Working replacement for ``eks/block_tridiag.py`` (whose ``Block_thomas_algo`` raises NameError
at :6): the block-tridiagonal Newton step of the EKS posterior, solved on the GPU.

The system solved is the consistent one: its solution from q = 0 is the exact
MAP path and equals the RTS smoothed mean; the diagonal blocks of H^-1 equal the RTS covariances.
"""
from __future__ import annotations

import numpy as np

from . import device as dev
from .ensemble_kalman import _cuda, _to_dev, _raise_if_singular, _params


def Block_thomas_algo(y, m0, S0, A, B, E, ensemble_vars, return_cov=False):
    """Signature of eks/block_tridiag.py:4.  y (T,p); ensemble_vars (T,p) [or the reference's (p,T)
    orientation when p != T, :16]; returns the Newton step s (T,r), optionally with the (T,r,r)
    diagonal blocks of H^-1."""
    d = _cuda()
    y = np.asarray(y, dtype=np.float64)
    ev = np.asarray(ensemble_vars, dtype=np.float64)
    T, p = y.shape
    if ev.shape == (p, T) and p != T:
        ev = ev.T
    prm = _params(d, m0, S0, B, A, E)
    obs = dev.pack_obs(_to_dev(y[None], d), _to_dev(ev[None], d))
    out = dev.btd_solve(obs, prm, 1, T)
    _raise_if_singular(out["status"])
    mean, cov = dev.expand_state(out["sol"], 1, T, prm)
    s = mean[0].cpu().numpy()
    return (s, cov[0].cpu().numpy()) if return_cov else s


def gradient_tridiag(y, m0, S0, A, B, E, ensemble_vars):
    """Gradient of the negative log posterior at q = 0, shape (r,T) like eks/block_tridiag.py:29-43,
    for the consistent system: G[:,t] = -B^T D_t^-1 y_t (and -S0^-1 m0 at t = 0).  Host numpy: it is
    an O(T) right-hand side used for inspection only; Block_thomas_algo builds it on the device."""
    y = np.asarray(y, dtype=np.float64)
    ev = np.asarray(ensemble_vars, dtype=np.float64)
    if ev.shape == (y.shape[1], y.shape[0]) and y.shape[0] != y.shape[1]:
        ev = ev.T
    G = -(np.asarray(B).T @ (y / ev).T)
    G[:, 0] -= np.linalg.solve(np.asarray(S0, dtype=np.float64), np.asarray(m0, dtype=np.float64))
    return G
