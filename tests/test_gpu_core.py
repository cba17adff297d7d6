"""GPU parity tests: the CUDA path (through the C ABI) against golden fixtures and the oracle."""
import numpy as np
import pandas as pd
import pytest

from conftest import load_golden, cases_of
from oracle import restated as o

pytestmark = pytest.mark.gpu

CORE = cases_of(load_golden("core_kats.npz"))
KF_CASES = [k for k in CORE if not k.startswith("ens_") and k != "denseAR"]
RTOL64 = 1e-5   # north_star: smoothed means/variances within 1e-5 relative in fp64 (we get ~1e-13)


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


@pytest.mark.parametrize("name", KF_CASES)
def test_filtering_pass_and_smoother_vs_reference(name):
    from eks_b200.ensemble_kalman import filtering_pass, smooth_backward
    from eks_b200.newton_eks import kalman_newton_recursive
    c = CORE[name]
    R = np.eye(c["C"].shape[0])
    mf, Vf, S = filtering_pass(c["y"], c["m0"], c["S0"], c["C"], R, c["A"], c["Q"], c["ev"])
    assert mf.dtype == np.float64 and Vf.shape == c["Vf"].shape and S.shape == c["S"].shape
    for a, b in ((mf, c["mf"]), (Vf, c["Vf"]), (S, c["S"])):
        assert rel(a, b) < 1e-11
    np.testing.assert_array_equal(R, c["R_after"])          # in-place diag(R) side effect
    np.testing.assert_array_equal(S[-1], c["S"][-1])         # S[T-1]==0 (or S0 when T==1)
    if "ms" in c:
        ms, Vs, CV = smooth_backward(c["y"], mf, Vf, S, c["A"], c["Q"], c["C"])
        assert CV.shape == c["CV"].shape
        for a, b in ((ms, c["ms"]), (Vs, c["Vs"]), (CV, c["CV"])):
            assert rel(a, b) < 1e-11
        q = kalman_newton_recursive(c["y"], c["m0"], c["S0"], c["A"], c["C"], c["ev"], c["Q"])
        assert rel(q, c["newton"]) < 1e-10


def test_dense_A_and_dense_R_generic_path():
    from eks_b200.ensemble_kalman import filtering_pass, smooth_backward
    c = CORE["denseAR"]
    R = c["R_in"].copy()
    mf, Vf, S = filtering_pass(c["y"], c["m0"], c["S0"], c["C"], R, c["A"], c["Q"], c["ev"])
    for a, b in ((mf, c["mf"]), (Vf, c["Vf"]), (S, c["S"]), (R, c["R_after"])):
        assert rel(a, b) < 1e-11
    ms, Vs, CV = smooth_backward(c["y"], mf, Vf, S, c["A"], c["Q"], c["C"])
    for a, b in ((ms, c["ms"]), (Vs, c["Vs"]), (CV, c["CV"])):
        assert rel(a, b) < 1e-10


def test_kalman_dot():
    from eks_b200.ensemble_kalman import kalman_dot
    rng = np.random.default_rng(5)
    for p, n in ((4, 3), (8, 3), (6, 4)):
        C = rng.normal(size=(p, n)); V = np.eye(n) + 0.1; R = np.diag(rng.uniform(0.1, 2, p)) + 0.05
        v = rng.normal(size=p); M = rng.normal(size=(p, n))
        np.testing.assert_allclose(kalman_dot(v, V, C, R), o.kalman_dot(v, V, C, R), rtol=1e-11, atol=1e-13)
        np.testing.assert_allclose(kalman_dot(M, V, C, R), o.kalman_dot(M, V, C, R), rtol=1e-11, atol=1e-13)
    with pytest.raises(np.linalg.LinAlgError):
        kalman_dot(np.ones(2), np.zeros((2, 2)), np.eye(2), np.zeros((2, 2)))


@pytest.mark.parametrize("name", [k for k in CORE if k.startswith("ens_")])
@pytest.mark.parametrize("mode", ["median", "mean"])
def test_ensemble_bit_exact(name, mode):
    from eks_b200.ensemble_kalman import ensemble
    c = CORE[name]
    st = c["stack"]
    keys = [f"k{i}" for i in range(st.shape[2])]
    dfs = [pd.DataFrame(st[m], columns=keys) for m in range(st.shape[0])]
    preds, evars, stacks, avg_d, var_d, stack_d = ensemble(dfs, keys, mode)
    np.testing.assert_array_equal(preds, c[f"{mode}_preds"])            # bit-exact, NaN positions included
    np.testing.assert_array_equal(evars, c[f"{mode}_vars"])             # same summation order -> bit-exact too
    np.testing.assert_array_equal(stacks, c["stacks_out"])
    assert stacks.strides == (8, 8 * st.shape[0], 8 * st.shape[0] * st.shape[1])
    np.testing.assert_array_equal(avg_d[keys[-1]], preds[:, -1])
    np.testing.assert_array_equal(stack_d[0][keys[0]], st[0, :, 0])


def test_ensemble_bad_mode():
    from eks_b200.ensemble_kalman import ensemble
    with pytest.raises(ValueError, match="averaging not supported"):
        ensemble([pd.DataFrame({"a": [1.0]})], ["a"], "mode")
