"""CPU: the per-thread programs of the CUDA kernels (eks_b200/csrc/eksb_math.cuh), compiled
for the host with g++ (tests/hostsim), against reference-generated goldens and the oracle.
This checks the information-form algebra the kernels use without needing a GPU."""
import numpy as np
import pytest

import hostsim
from conftest import load_golden, cases_of, rel
from oracle import restated as o

CORE = cases_of(load_golden("core_kats.npz"))
KF_CASES = [k for k in CORE if not k.startswith("ens_") and k != "denseAR"]


@pytest.mark.parametrize("name", KF_CASES)
@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-11), (np.float32, 1e-3)])
def test_kernel_math_vs_reference(name, dtype, tol):
    c = CORE[name]
    r = hostsim.kf_rts(c["y"], c["ev"], c["m0"], c["S0"], c["C"], c["A"], c["Q"], dtype=dtype, smooth="ms" in c)
    assert r["rc"] == 0
    assert rel(r["mf"], c["mf"]) < tol and rel(r["Vf"], c["Vf"]) < tol
    if "ms" in c:
        assert rel(r["ms"], c["ms"]) < tol and rel(r["Vs"], c["Vs"]) < tol and rel(r["CV"], c["CV"]) < tol
        assert rel(r["ym"], c["ms"] @ c["C"].T) < tol
        yv = np.einsum("jk,tkl,jl->tj", c["C"], c["Vs"], c["C"])
        assert rel(r["yv"], yv) < tol
    if dtype == np.float64:
        T, p = c["y"].shape
        want = o.nll(c["y"], c["m0"], c["S0"], c["C"], np.eye(p), c["A"], c["Q"], c["ev"])
        got = 0.5 * (r["nll_terms"] + T * p * np.log(2 * np.pi))
        assert abs(got - want) < 1e-9 * abs(want)


@pytest.mark.parametrize("name", ["p4", "p6", "p8", "n2p4", "T1", "T2"])
@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-11), (np.float32, 1e-4)])
def test_nll_split_form_vs_oracle(name, dtype, tol):
    """NLL from the helper record (Lam, iv, yy, log prod r) and the scaled chain == the oracle's Gaussian NLL."""
    c = CORE[name]
    T, p = c["y"].shape
    want = o.nll(c["y"], c["m0"], c["S0"], c["C"], np.eye(p), c["A"], c["Q"], c["ev"])
    r = hostsim.nll_split(c["y"], c["ev"], c["m0"], c["S0"], c["C"], c["A"], c["Q"], dtype=dtype)
    assert r["rc"] == 0 and abs(r["nll"] - want) < tol * abs(want)


def test_newton_first_frame_mode():
    c = CORE["p4"]
    S0i = np.linalg.inv(c["S0"])      # the kernel inverts S0 itself; the host build takes P0 directly
    r = hostsim.kf_rts(c["y"], c["ev"], c["m0"], S0i, c["C"], c["A"], c["Q"], first=2, smooth=False)
    assert rel(r["mf"], c["newton"]) < 1e-11


@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-9), (np.float32, 1e-3)])
def test_real_data_dynamic_range(dtype, tol):
    """mirror-mouse ensemble variances span 4e-4 .. 1.4e3 (SURVEY probe 4): the worst real
    conditioning the information form meets; fp32 must still meet north_star's 1e-3."""
    from oracle import restated_smoothers as osm
    z = cases_of(load_golden("mirror_mouse.npz"))
    for kp, c in z.items():
        st = osm.multicam_setup(c["stack"], 0.01, 25)
        R = np.eye(4)
        mf, Vf, S = o.filtering_pass(st["y"], st["m0"], st["S0"], st["C"], R, st["A"], st["Q"], st["ev"])
        ms, Vs, _ = o.smooth_backward(st["y"], mf, Vf, S, st["A"], st["Q"], st["C"])
        r = hostsim.kf_rts(st["y"], st["ev"], st["m0"], st["S0"], st["C"], st["A"], st["Q"], dtype=dtype, want_cv=False)
        assert r["rc"] == 0
        assert rel(r["ms"], ms) < tol and rel(r["Vs"], Vs) < tol and rel(r["ym"], ms @ st["C"].T) < tol


@pytest.mark.parametrize("name", [k for k in KF_CASES if "ms" in CORE[k]])
@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-11), (np.float32, 1e-3)])
def test_split_form_vs_reference(name, dtype, tol):
    """prep/chain split used by the warp-specialised kernels == reference."""
    c = CORE[name]
    r = hostsim.kf_rts_split(c["y"], c["ev"], c["m0"], c["S0"], c["C"], c["A"], c["Q"], dtype=dtype)
    assert r["rc"] == 0
    for k in ("mf", "Vf", "ms", "Vs"):
        assert rel(r[k], c[k]) < tol, k


@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-9), (np.float32, 1e-3)])
def test_split_form_real_data(dtype, tol):
    from oracle import restated_smoothers as osm
    z = cases_of(load_golden("mirror_mouse.npz"))
    for kp, c in z.items():
        st = osm.multicam_setup(c["stack"], 0.01, 25)
        mf, Vf, S = o.filtering_pass(st["y"], st["m0"], st["S0"], st["C"], np.eye(4), st["A"], st["Q"], st["ev"])
        ms, Vs, _ = o.smooth_backward(st["y"], mf, Vf, S, st["A"], st["Q"], st["C"])
        r = hostsim.kf_rts_split(st["y"], st["ev"], st["m0"], st["S0"], st["C"], st["A"], st["Q"], dtype=dtype)
        assert rel(r["ms"], ms) < tol and rel(r["Vs"], Vs) < tol and rel(r["mf"], mf) < tol


@pytest.mark.parametrize("name", [k for k in KF_CASES if "ms" in CORE[k]] + ["T1"])
@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-9), (np.float32, 1e-3)])
def test_block_tridiagonal_solve_equals_rts(name, dtype, tol):
    """A7: one Newton step of the block-tridiagonal system == RTS mean, H^-1 diagonal blocks == Vs
    (SURVEY fact 0.8), and == the oracle's block-Thomas restatement."""
    c = CORE[name]
    r = hostsim.btd(c["y"], c["ev"], c["m0"], c["S0"], c["C"], c["A"], c["Q"], dtype=dtype)
    assert r["rc"] == 0
    if "ms" in c:
        assert rel(r["x"], c["ms"]) < tol and rel(r["Sig"], c["Vs"]) < tol
    else:
        assert rel(r["x"], c["mf"]) < tol and rel(r["Sig"], c["Vf"]) < tol     # T == 1: smoother == filter
    x = o.btd_solve(c["y"], c["m0"], c["S0"], c["A"], c["C"], c["Q"], c["ev"])
    assert rel(r["x"], x) < tol


@pytest.mark.parametrize("name,chunk", [("p4", 16), ("p4", 257), ("p4", 300), ("p8", 7), ("n2p4", 5), ("n4p6", 9), ("T2", 1)])
@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-9), (np.float32, 1e-3)])
def test_chunked_scan_equals_sequential(name, chunk, dtype, tol):
    """K4: summaries + fold + re-run reproduce the sequential filter and smoother exactly
    (any chunk length, including ragged last chunks and a single chunk)."""
    c = CORE[name]
    r = hostsim.scan(c["y"], c["ev"], c["m0"], c["S0"], c["C"], c["A"], c["Q"], chunk, dtype=dtype)
    assert r["rc"] == 0
    for k in ("mf", "Vf", "ms", "Vs"):
        assert rel(r[k], c[k]) < tol, k


def _long_track(T, q, seed=0, p=4, n=3):
    rng = np.random.default_rng(seed)
    C = np.linalg.qr(rng.normal(size=(p, n)))[0]
    x = np.cumsum(rng.normal(size=(T, n)) * np.sqrt(q), axis=0) + 300.0        # far from 0: the cut must not care
    ev = np.exp(rng.normal(size=(T, p)) * 0.8 - 1.5)
    y = x @ C.T + np.sqrt(5 * ev) * rng.normal(size=(T, p))
    return y, ev, np.zeros(n), np.eye(n), C, np.eye(n), q * np.eye(n)


@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-10), (np.float32, 1e-3)])
def test_scan_early_writeout_equals_sequential(dtype, tol):
    """K4 early write-out: once a chunk's summary has forgotten the boundary state the summary pass
    writes the records itself; the result still equals the sequential recursion (and the scan without
    early write-out), and the cuts show that most of every chunk was written early."""
    T, L = 12000, 3000
    args = _long_track(T, 1e-2)
    seq = hostsim.kf_rts(*args, dtype=dtype, want_cv=False)
    r1 = hostsim.scan(*args, L, dtype=dtype, early=True)
    r0 = hostsim.scan(*args, L, dtype=dtype, early=False)
    assert r1["rc"] == 0 and r0["rc"] == 0
    for k in ("mf", "Vf", "ms", "Vs"):
        assert rel(r1[k], seq[k]) < tol, k
        assert rel(r1[k], r0[k]) < tol, k
    assert (r0["fcut"][1:] == L).all() and (r0["bcut"] == 0).all()            # nothing written early
    assert (r1["fcut"][1:] < L // 4).all() and (r1["fcut"][1:] > 0).all()     # re-run covers the warm-up only
    assert (r1["bcut"][:-1] > 3 * L // 4).all()


def test_scan_early_writeout_never_cuts_a_slowly_forgetting_filter():
    """With a tiny process noise the filter does not forget its start inside a chunk: no early write-out,
    full re-run, same result as the sequential recursion."""
    T, L = 6000, 1500
    args = _long_track(T, 1e-7, seed=1)
    seq = hostsim.kf_rts(*args, want_cv=False)
    r = hostsim.scan(*args, L, early=True)
    assert (r["fcut"][1:] == L).all() and (r["bcut"] == 0).all()
    for k in ("mf", "Vf", "ms", "Vs"):
        assert rel(r[k], seq[k]) < 1e-9, k


@pytest.mark.parametrize("dtype,tol", [(np.float64, 1e-9), (np.float32, 1e-3)])
def test_scan_mixed_chunks_some_forget_some_do_not(dtype, tol):
    """A stretch of uninformative frames (huge ensemble variance) keeps some chunks from forgetting their
    start: those go through the fold and a full re-run, their neighbours start from the records the summary
    pass wrote.  Still the sequential result."""
    T, L = 15000, 3000
    y, ev, m0, S0, C, A, Q = _long_track(T, 1e-2, seed=2)
    ev = ev.copy(); ev[5900:9100] *= 1e12
    seq = hostsim.kf_rts(y, ev, m0, S0, C, A, Q, dtype=dtype, want_cv=False)
    r = hostsim.scan(y, ev, m0, S0, C, A, Q, L, dtype=dtype, early=True)
    assert r["rc"] == 0
    assert r["fcut"][2] == L and r["bcut"][2] == 0                       # chunk 2 lies inside the blind stretch
    assert r["fcut"][4] < L and r["bcut"][0] > 0                         # far from it the early write-out works
    for k in ("mf", "Vf", "ms", "Vs"):
        assert rel(r[k], seq[k]) < tol, k
