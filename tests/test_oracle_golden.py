"""CPU: the oracle (oracle/restated*.py) against fixtures produced by the real reference."""
import numpy as np
import pandas as pd
import pytest

from conftest import load_golden, cases_of
from oracle import restated as o
from oracle import restated_smoothers as osm

CORE = cases_of(load_golden("core_kats.npz"))


@pytest.mark.parametrize("name", [k for k in CORE if not k.startswith("ens_") and k != "denseAR"])
def test_core_filter_smoother(name):
    c = CORE[name]
    R = np.eye(c["C"].shape[0])
    mf, Vf, S = o.filtering_pass(c["y"], c["m0"], c["S0"], c["C"], R, c["A"], c["Q"], c["ev"])
    for a, b in ((mf, c["mf"]), (Vf, c["Vf"]), (S, c["S"]), (R, c["R_after"])):
        np.testing.assert_allclose(a, b, rtol=1e-10, atol=1e-12)
    if "ms" in c:
        ms, Vs, CV = o.smooth_backward(c["y"], mf, Vf, S, c["A"], c["Q"], c["C"])
        for a, b in ((ms, c["ms"]), (Vs, c["Vs"]), (CV, c["CV"])):
            np.testing.assert_allclose(a, b, rtol=1e-10, atol=1e-12)
        q = o.kalman_newton_recursive(c["y"], c["m0"], c["S0"], c["A"], c["C"], c["ev"], c["Q"])
        np.testing.assert_allclose(q, c["newton"], rtol=1e-9, atol=1e-11)
        # block-tridiagonal identity: one Newton step from 0 == RTS mean
        x = o.btd_solve(c["y"], c["m0"], c["S0"], c["A"], c["C"], c["Q"], c["ev"])
        np.testing.assert_allclose(x, c["ms"], rtol=1e-8, atol=1e-10)


def test_core_dense_A_R():
    c = CORE["denseAR"]
    R = c["R_in"].copy()
    mf, Vf, S = o.filtering_pass(c["y"], c["m0"], c["S0"], c["C"], R, c["A"], c["Q"], c["ev"])
    ms, Vs, CV = o.smooth_backward(c["y"], mf, Vf, S, c["A"], c["Q"], c["C"])
    for a, b in ((mf, c["mf"]), (Vf, c["Vf"]), (S, c["S"]), (R, c["R_after"]), (ms, c["ms"]), (Vs, c["Vs"]), (CV, c["CV"])):
        np.testing.assert_allclose(a, b, rtol=1e-10, atol=1e-12)


@pytest.mark.parametrize("name", [k for k in CORE if k.startswith("ens_")])
@pytest.mark.parametrize("mode", ["median", "mean"])
def test_ensemble(name, mode):
    c = CORE[name]
    st = c["stack"]
    pr, va = o.ensemble_arrays(st, mode)
    np.testing.assert_array_equal(pr, c[f"{mode}_preds"])          # bit-exact incl. NaN positions
    np.testing.assert_allclose(va, c[f"{mode}_vars"], rtol=1e-13, atol=0, equal_nan=True)
    keys = [f"k{i}" for i in range(st.shape[2])]
    dfs = [pd.DataFrame(st[m], columns=keys) for m in range(st.shape[0])]
    out = o.ensemble(dfs, keys, mode)
    np.testing.assert_array_equal(out[0], c[f"{mode}_preds"])
    np.testing.assert_array_equal(out[2], c["stacks_out"])
    assert out[2].shape == st.shape and out[2].strides == (8, 8 * st.shape[0], 8 * st.shape[0] * st.shape[1])
    assert set(out[5].keys()) == set(range(st.shape[0]))


def test_ensemble_bad_mode():
    with pytest.raises(ValueError, match="averaging not supported"):
        o.ensemble_arrays(np.zeros((2, 3, 1)), "mode")


def _cams(stack, names):
    M = stack.shape[0]
    return [[pd.DataFrame(stack[m][:, 2 * c:2 * c + 2], columns=["kp_x", "kp_y"]) for m in range(M)]
            for c in range(len(names))]


def test_mirror_mouse():
    z = cases_of(load_golden("mirror_mouse.npz"))
    for kp, c in z.items():
        cams = ["top", "bot"]
        std = osm.ensemble_kalman_smoother_multi_cam(_cams(c["stack"], cams), "kp", 0.01, 25, cams)
        got = np.column_stack([std[f"{n}_df"].to_numpy()[:, :2] for n in cams])
        np.testing.assert_allclose(got, c["ref_std"], rtol=0, atol=1e-9)
        np.testing.assert_allclose(got, c["shipped_std"], rtol=0, atol=1e-9)
        assert std["top_df"].columns.names == ["scorer", "bodyparts", "coords"]
        assert np.isnan(std["top_df"].to_numpy()[:, 2]).all()
        opt = osm.eks_opti_smoother_multi_cam(_cams(c["stack"], cams), "kp", 0.01, 25, cams)
        got = np.column_stack([opt[f"{n}_df"].to_numpy()[:, :2] for n in cams])
        np.testing.assert_allclose(got, c["ref_opti"], rtol=0, atol=1e-9)
        np.testing.assert_allclose(got, c["shipped_opti"], rtol=0, atol=1e-9)


def test_fish():
    z = cases_of(load_golden("fish.npz"))
    cams = ["main", "top", "right"]
    assert len(z) >= 50
    for kp, c in z.items():
        out = osm.ensemble_kalman_smoother_multi_cam(_cams(c["stack"], cams), "kp", 0.01, 50, cams)
        got = np.column_stack([out[f"{n}_df"].to_numpy()[:, :2] for n in cams])
        np.testing.assert_allclose(got, c["ref"], rtol=0, atol=1e-8)
        np.testing.assert_allclose(got, c["shipped"], rtol=0, atol=1e-8)


def test_pupil():
    z = load_golden("pupil.npz")
    dfs = [pd.DataFrame(z["stack"][m], columns=osm.PUPIL_KEYS) for m in range(z["stack"].shape[0])]
    kp = [str(s) for s in z["keypoint_names"]]
    for tag, A in (("a99", np.diag([.99, .99, .99])), ("adef", np.diag([.9999, .999, .999]))):
        out = osm.ensemble_kalman_smoother_pupil(dfs, kp, "ensemble-kalman_tracker", A)
        np.testing.assert_allclose(out["latents_df"].to_numpy(), z[f"{tag}/latents"], rtol=0, atol=1e-10)
        np.testing.assert_allclose(out["markers_df"].to_numpy(), z[f"{tag}/markers"], rtol=0, atol=1e-10, equal_nan=True)
    np.testing.assert_allclose(z["a99/latents"], z["shipped/latents"], rtol=0, atol=1e-10)
    np.testing.assert_allclose(osm.eks_opti_pupil_latents(dfs), z["shipped/opti_latents"], rtol=0, atol=1e-10)


def test_paw():
    z = load_golden("paw.npz")
    cols = ["paw_l_x", "paw_l_y", "paw_l_likelihood", "paw_r_x", "paw_r_y", "paw_r_likelihood"]
    left = [pd.DataFrame(a, columns=cols) for a in z["left"]]
    right = [pd.DataFrame(a, columns=cols) for a in z["right"]]
    out = osm.ensemble_kalman_smoother_paw_asynchronous(
        left, right, z["ts_left"], z["ts_right"], [str(s) for s in z["keypoint_names"]], 0.01, 25)
    np.testing.assert_allclose(out["left_df"].to_numpy(), z["left_df"], rtol=0, atol=1e-9, equal_nan=True)
    np.testing.assert_allclose(out["right_df"].to_numpy(), z["right_df"], rtol=0, atol=1e-9, equal_nan=True)


def test_paw_opti_twin():
    z = load_golden("paw.npz")
    cols = ["paw_l_x", "paw_l_y", "paw_l_likelihood", "paw_r_x", "paw_r_y", "paw_r_likelihood"]
    left = [pd.DataFrame(a, columns=cols) for a in z["left"]]
    right = [pd.DataFrame(a, columns=cols) for a in z["right"]]
    out = osm.eks_opti_smoother_paw_asynchronous(
        left, right, z["ts_left"], z["ts_right"], [str(s) for s in z["keypoint_names"]], 0.01, 25)
    np.testing.assert_allclose(out["left_df"].to_numpy(), z["opti_left_df"], rtol=0, atol=1e-9, equal_nan=True)
    np.testing.assert_allclose(out["right_df"].to_numpy(), z["opti_right_df"], rtol=0, atol=1e-9, equal_nan=True)


def test_pupil_opti_markers_as_shipped():
    """data/misc/pupil-test/opti_eks_pupil_traces.csv: markers rebuilt from the standard smoother's latents with
    the means added twice (pupil_smoother.py:325-346)."""
    z = load_golden("pupil.npz")
    dfs = [pd.DataFrame(a, columns=osm.PUPIL_KEYS) for a in z["stack"]]
    kp = [str(s) for s in z["keypoint_names"]]
    out = osm.eks_opti_smoother_pupil(dfs, kp, "ensemble-kalman_tracker")
    np.testing.assert_allclose(out["markers_df"].to_numpy(), z["shipped/opti_markers"], rtol=0, atol=1e-9, equal_nan=True)
    np.testing.assert_allclose(out["latents_df"].to_numpy(), z["shipped/opti_latents"], rtol=0, atol=1e-10)


EXTRA = cases_of(load_golden("extra_kats.npz"))


@pytest.mark.parametrize("name", ["zerovar", "n5p10", "n5p4", "n7p12", "n3p10"])
def test_oracle_extra_shapes_and_zero_variance(name):
    c = EXTRA[name]
    R = np.eye(c["C"].shape[0])
    mf, Vf, S = o.filtering_pass(c["y"], c["m0"], c["S0"], c["C"], R, c["A"], c["Q"], c["ev"])
    ms, Vs, CV = o.smooth_backward(c["y"], mf, Vf, S, c["A"], c["Q"], c["C"])
    for a, b in ((mf, c["mf"]), (Vf, c["Vf"]), (S, c["S"]), (ms, c["ms"]), (Vs, c["Vs"]), (CV, c["CV"])):
        assert np.isfinite(b).all()
        np.testing.assert_allclose(a, b, rtol=0, atol=1e-10 * max(1.0, np.abs(b).max()))
    if "newton" in c:
        np.testing.assert_allclose(o.kalman_newton_recursive(c["y"], c["m0"], c["S0"], c["A"], c["C"], c["ev"], c["Q"]),
                                   c["newton"], rtol=0, atol=1e-10)


def test_oracle_newton_sweeps_and_own_S():
    c = EXTRA["newton3"]
    q3, loss = o.kalman_newton_recursive(c["y"], c["m0"], c["S0"], c["A"], c["C"], c["ev"], c["Q"], max_iter=3)
    np.testing.assert_allclose(q3, c["q3"], rtol=0, atol=1e-10)
    np.testing.assert_array_equal(loss, c["loss3"])
    assert np.abs(c["q3"] - c["q1"]).max() > 1e-6              # the extra sweeps do change q (P is carried on)
    c = EXTRA["ownS"]
    ms, Vs, CV = o.smooth_backward(c["y"], c["mf"], c["Vf"], c["S_own"], c["A"], c["Q"], c["C"])
    for a, b in ((ms, c["ms"]), (Vs, c["Vs"]), (CV, c["CV"])):
        np.testing.assert_allclose(a, b, rtol=0, atol=1e-10)


def test_oracle_twenty_model_ensemble_and_five_cameras():
    c = EXTRA["ens_m20"]
    for mode in ("median", "mean"):
        pr, va = o.ensemble_arrays(c["stack"], mode)
        np.testing.assert_array_equal(pr, c[f"{mode}_preds"])
        np.testing.assert_array_equal(va, c[f"{mode}_vars"])
    c = EXTRA["multicam5"]
    st = osm.multicam_setup(c["stack"], 0.01, 25)
    mf, Vf, S = o.filtering_pass(st["y"], st["m0"], st["S0"], st["C"], np.eye(10), st["A"], st["Q"], st["ev"])
    ms, _, _ = o.smooth_backward(st["y"], mf, Vf, S, st["A"], st["Q"], st["C"])
    np.testing.assert_allclose(ms @ st["C"].T + st["means"], c["ref"], rtol=0, atol=1e-8)


def test_nll_matches_dense_gaussian():
    from scipy.stats import multivariate_normal
    rng = np.random.default_rng(3)
    T, p, n = 5, 4, 3
    C = rng.normal(size=(p, n)); A = np.diag([.9, .95, 1.0]); Q = 0.1 * np.eye(n) + 0.02
    S0 = np.diag([1.0, 2.0, 0.5]); m0 = rng.normal(size=n)
    y = rng.normal(size=(T, p)); ev = np.exp(rng.normal(size=(T, p)))
    P = [S0]; mu = [m0]
    for t in range(1, T):
        P.append(A @ P[-1] @ A.T + Q); mu.append(A @ mu[-1])
    cov = np.zeros((T * n, T * n))
    for i in range(T):
        for j in range(i, T):
            blk = P[i] @ np.linalg.matrix_power(A.T, j - i)
            cov[i*n:(i+1)*n, j*n:(j+1)*n] = blk
            cov[j*n:(j+1)*n, i*n:(i+1)*n] = blk.T
    Cb = np.kron(np.eye(T), C)
    want = -multivariate_normal.logpdf(y.reshape(-1), Cb @ np.concatenate(mu), Cb @ cov @ Cb.T + np.diag(ev.reshape(-1)))
    assert abs(o.nll(y, m0, S0, C, np.eye(p), A, Q, ev) - want) < 1e-9


def test_btd_dense_blocks_equal_rts_cov():
    rng = np.random.default_rng(4)
    T, p, n = 12, 4, 3
    C = np.linalg.qr(rng.normal(size=(p, n)))[0]; A = np.diag([.9, .95, 1.0]); Q = 0.1 * np.eye(n)
    S0 = np.eye(n); m0 = rng.normal(size=n); y = rng.normal(size=(T, p)); ev = np.exp(rng.normal(size=(T, p)))
    mf, Vf, S = o.filtering_pass(y, m0, S0, C, np.eye(p), A, Q, ev)
    ms, Vs, CV = o.smooth_backward(y, mf, Vf, S, A, Q, C)
    x, Hi = o.btd_solve_dense(y, m0, S0, A, C, Q, ev)
    np.testing.assert_allclose(x, ms, atol=1e-12)
    for t in range(T):
        np.testing.assert_allclose(Hi[t*n:(t+1)*n, t*n:(t+1)*n], Vs[t], atol=1e-12)
    np.testing.assert_allclose(o.btd_solve(y, m0, S0, A, C, Q, ev), ms, atol=1e-12)
