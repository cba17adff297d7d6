"""ctypes access to the host build of the kernel math (test infrastructure only)."""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libhostsim.so")


def build(force=False):
    srcs = [os.path.join(HERE, "hostsim.cpp"),
            os.path.join(HERE, "../../eks_b200/csrc/eksb_math.cuh"),
            os.path.join(HERE, "../../eks_b200/csrc/eksb_dims.h")]
    if force or not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in srcs):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-ffp-contract=off",
                               "-x", "c++", srcs[0], "-o", LIB])
    return ctypes.CDLL(LIB)


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = build()
    return _lib


def sym_unpack(packed, n):
    """(T, n(n+1)/2) packed upper-triangular rows -> (T,n,n)."""
    T = packed.shape[0]
    out = np.zeros((T, n, n), dtype=packed.dtype)
    k = 0
    for i in range(n):
        for j in range(i, n):
            out[:, i, j] = packed[:, k]
            out[:, j, i] = packed[:, k]
            k += 1
    return out


def a_mode(A):
    return 0 if np.count_nonzero(A - np.diag(np.diagonal(A))) == 0 else 1


def kf_rts(y, ev, m0, S0, C, A, Q, dtype=np.float64, first=1, smooth=True, want_cv=True):
    """Run the host build of kf_forward_range + rts_backward_range on one track."""
    T, p = y.shape
    n = m0.shape[0]
    c = lambda a: np.ascontiguousarray(a, dtype=dtype)
    am = a_mode(A)
    y, ev, m0, S0, C, A, Q = map(c, (y, ev, m0, S0, C, A, Q))
    sg = n + n * (n + 1) // 2
    filt = np.zeros((T, sg), dtype=dtype)
    smth = np.zeros((T, sg), dtype=dtype)
    ym = np.zeros((T, p), dtype=dtype)
    yv = np.zeros((T, p), dtype=dtype)
    cv = np.zeros((T, n * n), dtype=dtype)
    nll = ctypes.c_double(0.0)
    fn = lib().hs_kf_rts_f64 if dtype == np.float64 else lib().hs_kf_rts_f32
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    null = ctypes.c_void_p(0)
    rc = fn(ctypes.c_int(n), ctypes.c_int(p), ctypes.c_int(am), ctypes.c_long(T), P(y), P(ev), P(m0), P(S0), P(C),
            P(A), P(Q), ctypes.c_int(first), P(filt), P(smth) if smooth else null, P(ym), P(yv),
            P(cv) if want_cv else null, ctypes.byref(nll))
    if rc < 0:
        raise ValueError(f"unsupported dims n={n} p={p}")
    out = dict(rc=rc, mf=filt[:, :n], Vf=sym_unpack(filt[:, n:], n), nll_terms=nll.value)
    if smooth:
        out.update(ms=smth[:, :n], Vs=sym_unpack(smth[:, n:], n), ym=ym, yv=yv, CV=cv.reshape(T, n, n)[:T - 1])
    return out


def kf_rts_split(y, ev, m0, S0, C, A, Q, dtype=np.float64, first=1):
    """Host build of the prep/chain split (fwd_prep+fwd_chain, bwd_prep+bwd_chain) on one track."""
    T, p = y.shape
    n = m0.shape[0]
    c = lambda a: np.ascontiguousarray(a, dtype=dtype)
    am = a_mode(A)
    y, ev, m0, S0, C, A, Q = map(c, (y, ev, m0, S0, C, A, Q))
    sg = n + n * (n + 1) // 2
    filt = np.zeros((T, sg), dtype=dtype)
    smth = np.zeros((T, sg), dtype=dtype)
    fn = lib().hs_split_f64 if dtype == np.float64 else lib().hs_split_f32
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    rc = fn(ctypes.c_int(n), ctypes.c_int(p), ctypes.c_int(am), ctypes.c_long(T), P(y), P(ev), P(m0), P(S0), P(C),
            P(A), P(Q), ctypes.c_int(first), P(filt), P(smth))
    if rc < 0:
        raise ValueError(f"unsupported dims n={n} p={p}")
    return dict(rc=rc, mf=filt[:, :n], Vf=sym_unpack(filt[:, n:], n), ms=smth[:, :n], Vs=sym_unpack(smth[:, n:], n))


def nll_split(y, ev, m0, S0, C, A, Q, dtype=np.float64, first=1):
    """Host build of the NLL in the prep / chain split (fwd_prep_outer_nll + fwd_nll_scaled); diagonal A, n <= 3.
    Returns 0.5 * (sum of per-frame terms + T p log 2 pi), like the kernels."""
    T, p = y.shape
    n = m0.shape[0]
    c = lambda a: np.ascontiguousarray(a, dtype=dtype)
    assert a_mode(A) == 0
    y, ev, m0, S0, C, A, Q = map(c, (y, ev, m0, S0, C, A, Q))
    out = ctypes.c_double(0.0)
    fn = lib().hs_nll_split_f64 if dtype == np.float64 else lib().hs_nll_split_f32
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    rc = fn(ctypes.c_int(n), ctypes.c_int(p), ctypes.c_long(T), P(y), P(ev), P(m0), P(S0), P(C), P(A), P(Q),
            ctypes.c_int(first), ctypes.byref(out))
    if rc < 0:
        raise ValueError(f"unsupported dims n={n} p={p}")
    return dict(rc=rc, nll=0.5 * (out.value + T * p * np.log(2 * np.pi)))


def btd(y, ev, m0, S0, C, A, Q, dtype=np.float64):
    """Host build of the block-tridiagonal solve (fwd_prep + btd_fwd_chain, btd_bwd_prep + bwd_chain)."""
    T, p = y.shape
    n = m0.shape[0]
    c = lambda a: np.ascontiguousarray(a, dtype=dtype)
    am = a_mode(A)
    y, ev, m0, S0, C, A, Q = map(c, (y, ev, m0, S0, C, A, Q))
    sg = n + n * (n + 1) // 2
    work = np.zeros((T, sg), dtype=dtype)
    sol = np.zeros((T, sg), dtype=dtype)
    fn = lib().hs_btd_f64 if dtype == np.float64 else lib().hs_btd_f32
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    rc = fn(ctypes.c_int(n), ctypes.c_int(p), ctypes.c_int(am), ctypes.c_long(T), P(y), P(ev), P(m0), P(S0), P(C),
            P(A), P(Q), P(work), P(sol))
    if rc < 0:
        raise ValueError(f"unsupported dims n={n} p={p}")
    return dict(rc=rc, x=sol[:, :n], Sig=sym_unpack(sol[:, n:], n))


def scan(y, ev, m0, S0, C, A, Q, chunk, dtype=np.float64, early=True):
    """Host run of the chunked parallel-in-time phases (summaries, fold, re-run) on one track.
    early: the summary pass writes the records itself once it has forgotten the chunk's boundary state
    (the device default); fcut / bcut in the result are the per-chunk cut positions."""
    T, p = y.shape
    n = m0.shape[0]
    c = lambda a: np.ascontiguousarray(a, dtype=dtype)
    am = a_mode(A)
    y, ev, m0, S0, C, A, Q = map(c, (y, ev, m0, S0, C, A, Q))
    sg = n + n * (n + 1) // 2
    filt = np.zeros((T, sg), dtype=dtype)
    smth = np.zeros((T, sg), dtype=dtype)
    fn = lib().hs_scan_f64 if dtype == np.float64 else lib().hs_scan_f32
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    nc = (T + chunk - 1) // chunk
    cuts = np.zeros(2 * nc, dtype=np.int64)
    rc = fn(ctypes.c_int(n), ctypes.c_int(p), ctypes.c_int(am), ctypes.c_long(T), ctypes.c_long(chunk),
            ctypes.c_int(1 if early else 0), P(y), P(ev), P(m0), P(S0), P(C), P(A), P(Q), P(filt), P(smth), P(cuts))
    if rc < 0:
        raise ValueError(f"unsupported dims n={n} p={p}")
    return dict(rc=rc, fcut=cuts[:nc], bcut=cuts[nc:], mf=filt[:, :n], Vf=sym_unpack(filt[:, n:], n), ms=smth[:, :n], Vs=sym_unpack(smth[:, n:], n))
