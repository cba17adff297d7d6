"""GPU: the reference-named smoothers against the reference's own shipped outputs (golden CSV values
frozen in tests/golden by make_golden.py) -- mirror-mouse, ibl-pupil, ibl-paw, mirror-fish clips."""
import numpy as np
import pandas as pd
import pytest

from conftest import load_golden, cases_of

pytestmark = pytest.mark.gpu
TOL = 1e-7     # absolute, pixels; observed ~1e-10.  north_star bar: 1e-5 relative in fp64


def _cams(stack, names):
    M = stack.shape[0]
    return [[pd.DataFrame(stack[m][:, 2 * c:2 * c + 2], columns=["kp_x", "kp_y"]) for m in range(M)]
            for c in range(len(names))]


def test_mirror_mouse_standard_and_opti():
    from eks_b200.multiview_pca_smoother import ensemble_kalman_smoother_multi_cam, eks_opti_smoother_multi_cam
    cams = ["top", "bot"]
    for kp, c in cases_of(load_golden("mirror_mouse.npz")).items():
        out = ensemble_kalman_smoother_multi_cam(_cams(c["stack"], cams), kp, 0.01, 25, cams)
        got = np.column_stack([out[f"{n}_df"].to_numpy()[:, :2] for n in cams])
        np.testing.assert_allclose(got, c["shipped_std"], rtol=0, atol=TOL)
        np.testing.assert_allclose(got, c["ref_std"], rtol=0, atol=TOL)
        df = out["top_df"]
        assert list(df.columns) == [("ensemble-kalman_tracker", kp, k) for k in ("x", "y", "likelihood")]
        assert df.columns.names == ["scorer", "bodyparts", "coords"] and np.isnan(df.to_numpy()[:, 2]).all()
        opt = eks_opti_smoother_multi_cam(_cams(c["stack"], cams), kp, 0.01, 25, cams, plot=False)
        got = np.column_stack([opt[f"{n}_df"].to_numpy()[:, :2] for n in cams])
        np.testing.assert_allclose(got, c["shipped_opti"], rtol=0, atol=TOL)


def test_fish_clips():
    from eks_b200.multiview_pca_smoother import ensemble_kalman_smoother_multi_cam
    cams = ["main", "top", "right"]
    z = cases_of(load_golden("fish.npz"))
    assert len(z) >= 100
    for kp, c in z.items():
        out = ensemble_kalman_smoother_multi_cam(_cams(c["stack"], cams), "kp", 0.01, 50, cams)
        got = np.column_stack([out[f"{n}_df"].to_numpy()[:, :2] for n in cams])
        np.testing.assert_allclose(got, c["shipped"], rtol=0, atol=1e-6)


def test_pupil_standard_and_opti():
    from eks_b200.pupil_smoother import ensemble_kalman_smoother_pupil, eks_opti_smoother_pupil, KEYS
    z = load_golden("pupil.npz")
    dfs = [pd.DataFrame(z["stack"][m], columns=KEYS) for m in range(z["stack"].shape[0])]
    kp = [str(s) for s in z["keypoint_names"]]
    out = ensemble_kalman_smoother_pupil(dfs, kp, "ensemble-kalman_tracker", np.diag([.99, .99, .99]))
    np.testing.assert_allclose(out["latents_df"].to_numpy(), z["shipped/latents"], rtol=0, atol=TOL)
    np.testing.assert_allclose(out["markers_df"].to_numpy(), z["shipped/markers"], rtol=0, atol=TOL, equal_nan=True)
    assert out["markers_df"].shape == (2000, 12) and out["latents_df"].columns.names == ["scorer", "latent"]
    assert [c[1] for c in out["latents_df"].columns] == ["diameter", "com_x", "com_y"]
    out = ensemble_kalman_smoother_pupil(dfs, kp, "ensemble-kalman_tracker", np.diag([.9999, .999, .999]))
    np.testing.assert_allclose(out["latents_df"].to_numpy(), z["adef/latents"], rtol=0, atol=TOL)
    opt = eks_opti_smoother_pupil(dfs, kp, "ensemble-kalman_tracker", np.eye(3), plot=False)
    np.testing.assert_allclose(opt["latents_df"].to_numpy(), z["shipped/opti_latents"], rtol=0, atol=TOL)


def test_paw_asynchronous():
    from eks_b200.multiview_pca_smoother import ensemble_kalman_smoother_paw_asynchronous, eks_opti_smoother_paw_asynchronous
    from oracle import restated as o
    z = load_golden("paw.npz")
    cols = ["paw_l_x", "paw_l_y", "paw_l_likelihood", "paw_r_x", "paw_r_y", "paw_r_likelihood"]
    left = [pd.DataFrame(a, columns=cols) for a in z["left"]]
    right = [pd.DataFrame(a, columns=cols) for a in z["right"]]
    kp = [str(s) for s in z["keypoint_names"]]
    out = ensemble_kalman_smoother_paw_asynchronous(left, right, z["ts_left"], z["ts_right"], kp, 0.01, 25)
    assert out["left_df"].shape == (997, 6)
    np.testing.assert_allclose(out["left_df"].to_numpy(), z["left_df"], rtol=0, atol=TOL, equal_nan=True)
    np.testing.assert_allclose(out["right_df"].to_numpy(), z["right_df"], rtol=0, atol=TOL, equal_nan=True)
    opt = eks_opti_smoother_paw_asynchronous(left, right, z["ts_left"], z["ts_right"], kp, 0.01, 25)
    assert opt["left_df"].shape == (997, 6) and np.isfinite(opt["left_df"].to_numpy()[:, [0, 1, 3, 4]]).all()


def _synthetic_session(B, T, p, M=5, seed=0):
    rng = np.random.default_rng(seed)
    C = np.stack([np.linalg.qr(rng.normal(size=(p, 3)))[0] for _ in range(B)])
    x = np.cumsum(rng.normal(size=(B, T, 3)) * 0.3, axis=1)
    clean = np.einsum("bpn,btn->btp", C, x) + rng.uniform(50, 400, size=(B, 1, p))
    sd = np.exp(rng.normal(np.log(0.5), 0.8, size=(B, T, p)))
    sd[rng.random((B, T, 1)).repeat(p, 2) < 0.02] *= 30
    return (clean[None] + sd[None] * rng.normal(size=(M, B, T, p))).transpose(0, 2, 1, 3)   # (M,T,B,p)


@pytest.mark.parametrize("B,T,p,q", [(5, 700, 4, 25), (37, 333, 6, 50), (3, 1500, 8, 90), (2, 64, 4, 100), (3, 40000, 4, 25)])
def test_device_parameter_setup_matches_oracle(B, T, p, q):
    """K8: percentile threshold / good-frame means / PCA / S0 / Q computed on the device equal the oracle's
    restatement of multiview_pca_smoother.py:684-731 (numpy percentile + sklearn PCA) per track.  (3 tracks x 40 000
    frames: 157 moment slices per tile, folded to 32 by setup_reduce_kernel before the finalize.)"""
    import torch
    from eks_b200 import device as dev
    from oracle import restated_smoothers as osm
    stack = _synthetic_session(B, T, p, seed=B + T)
    d = torch.device("cuda", 0)
    preds = torch.as_tensor(np.ascontiguousarray(stack.reshape(5, T, B * p))).to(d)
    obs = dev.ensemble_tiled(preds, B, p, "median")
    prm, means, status = dev.pca_setup(obs, B, T, p, q, 0.01, center=True)
    assert int(status.sum()) == 0
    y_c = dev.unpack_rows(obs, 2 * p, 0, B, T, p).cpu().numpy()
    for b in range(B):
        want = osm.multicam_setup(stack[:, :, b, :], 0.01, q)
        np.testing.assert_allclose(means[b].cpu().numpy(), want["means"], rtol=1e-12)
        np.testing.assert_allclose(y_c[b], want["y"], rtol=0, atol=1e-9)
        Cg, Cw = prm.C[b].cpu().numpy(), want["C"]
        sgn = np.sign(np.sum(Cg * Cw, axis=0))                       # component signs are a convention
        np.testing.assert_allclose(Cg * sgn, Cw, rtol=0, atol=1e-8)
        np.testing.assert_allclose(prm.S0[b].cpu().numpy(), want["S0"], rtol=1e-8, atol=1e-12)
        Qw = want["Q"] * np.outer(sgn, sgn)
        np.testing.assert_allclose(prm.Q[b].cpu().numpy(), Qw, rtol=1e-7, atol=1e-9 * np.abs(Qw).max())


@pytest.mark.parametrize("dtname,rtol", [("float64", 1e-10), ("float32", 2e-5)])
@pytest.mark.parametrize("case", ["plain", "ties", "many_ties", "one_bin", "tiny"])
def test_percentile_select_candidate_buffer_and_overflow_fallback(dtname, rtol, case, monkeypatch):
    """The exact percentile is a radix select.  Default: ONE full-length pass with a wide first digit (sign + exponent +
    3 mantissa bits), the keys of its bin collected into a short buffer, the remaining bits resolved there.  Same
    parameters, bit for bit, as the round-1 path (three 8-bit passes, then collect: EKSB_SETUP_NO_WIDE=1) and as every
    pass at full length (EKSB_SETUP_NO_COLLECT=1) -- also off the common case: more keys in the bin than candidate slots
    ("one_bin": 30 000 variances in [1, 1.05] > 8192 slots -> one more full-length 8-bit digit, then collect again),
    more TIED keys than slots ("many_ties": 10 000 copies -> the second collect overflows too, remaining passes at full
    length), or a quantile among variances below the wide digit's window ("tiny": a third of the frames at 1e-30 ->
    the 8-bit passes).  Good-frame means equal the numpy percentile mask."""
    import torch
    from eks_b200 import device as dev
    from oracle import restated_smoothers as osm
    dt = getattr(torch, dtname)
    B, p = 40, 4
    T = 30000 if case in ("many_ties", "one_bin") else 6000
    rng = np.random.default_rng(11 + len(case))
    y = np.cumsum(rng.normal(size=(B, T, p)) * 0.3, axis=1) + rng.uniform(50, 400, size=(B, 1, p))
    ev = np.exp(rng.normal(np.log(0.05), 1.6, size=(B, T, p)))
    if case in ("ties", "many_ties"):                                 # 3 distinct max_vars, T/3 copies of each
        ev = np.array([0.01, 0.1, 1.0])[rng.integers(0, 3, size=(B, T, 1))] * np.ones(p)
    if case == "one_bin":
        ev = 1.0 + 0.05 * rng.random((B, T, p))
    if case == "tiny":
        ev = np.where(rng.random((B, T, 1)) < 0.34, 1e-30 * (1 + rng.random((B, T, p))), ev)
    d = torch.device("cuda", 0)
    yt, evt = torch.as_tensor(y, dtype=dt), torch.as_tensor(ev, dtype=dt)
    obs = dev.pack_obs(yt.to(d), evt.to(d))
    for q in (25, 60):
        res = {}
        for mode in ("wide", "narrow", "full"):
            monkeypatch.delenv("EKSB_SETUP_NO_COLLECT", raising=False)
            monkeypatch.delenv("EKSB_SETUP_NO_WIDE", raising=False)
            if mode == "full":
                monkeypatch.setenv("EKSB_SETUP_NO_COLLECT", "1")
            if mode == "narrow":
                monkeypatch.setenv("EKSB_SETUP_NO_WIDE", "1")
            prm, means, status = dev.pca_setup(obs.clone(), B, T, p, q, 0.01, center=False)
            assert int(status.sum()) == 0
            res[mode] = (means, prm.C, prm.S0, prm.Q)
        for mode in ("wide", "narrow"):
            for a, b in zip(res[mode], res["full"]):
                assert torch.equal(a, b), (mode, q, float((a.double() - b.double()).abs().max()))
        for b in (0, 17, B - 1):
            good = osm.good_frame_mask(evt[b].double().numpy(), q)
            np.testing.assert_allclose(res["wide"][0][b].cpu().double().numpy(), yt[b].double().numpy()[good].mean(axis=0), rtol=rtol)


@pytest.mark.parametrize("dtname", ["float64", "float32"])
@pytest.mark.parametrize("B,T,p", [(37, 333, 6), (70, 900, 4)])
def test_max_vars_from_the_ensemble_kernel_equal_the_setup_pass(dtname, B, T, p):
    """eksb_ensemble_tiled_mx writes max_j ev[t,j] into the setup workspace while the variances are in registers;
    eksb_pca_setup_mx(mx_ready=1) then skips its own pass: same records, same parameters, bit for bit."""
    import torch
    from eks_b200 import device as dev
    dt = getattr(torch, dtname)
    stack = _synthetic_session(B, T, p, seed=3 * B + T)
    d = torch.device("cuda", 0)
    preds = torch.as_tensor(np.ascontiguousarray(stack.reshape(5, T, B * p)), dtype=dt).to(d)
    obs0 = dev.ensemble_tiled(preds, B, p, "median")
    ref = dev.pca_setup(obs0.clone(), B, T, p, 25, 0.01, center=False)
    ws = dev.setup_workspace(B, T, p, dt, d)
    obs1 = dev.ensemble_tiled(preds, B, p, "median", max_var=dev.setup_max_var_view(ws, B, T, p, dt))
    assert torch.equal(obs0, obs1)
    got = dev.pca_setup(obs1, B, T, p, 25, 0.01, center=False, workspace=ws, mx_ready=True)
    assert int(got[2].sum()) == 0
    assert torch.equal(got[1], ref[1])
    for a, b in ((got[0].C, ref[0].C), (got[0].S0, ref[0].S0), (got[0].Q, ref[0].Q)):
        assert torch.equal(a, b)


def test_session_entry_equals_per_keypoint_calls():
    """All mirror-mouse keypoints in one batched session call == the per-keypoint drop-in == shipped CSVs."""
    from eks_b200.multiview_pca_smoother import smooth_session_multi_cam, session_dfs
    z = cases_of(load_golden("mirror_mouse.npz"))
    kps = sorted(z)
    stack = np.stack([z[k]["stack"] for k in kps], axis=2)           # (M,T,B,p)
    res = smooth_session_multi_cam(stack, 0.01, 25)
    for i, k in enumerate(kps):
        np.testing.assert_allclose(res["y"][i], z[k]["shipped_std"], rtol=0, atol=TOL)
    dfs = session_dfs(res["y"], kps, ["top", "bot"])
    assert dfs["top_df"].shape == (stack.shape[1], 3 * len(kps))
    assert list(dfs["bot_df"].columns[:3]) == [("ensemble-kalman_tracker", kps[0], c) for c in ("x", "y", "likelihood")]
    opt = smooth_session_multi_cam(stack, 0.01, 25, newton=True)
    for i, k in enumerate(kps):
        np.testing.assert_allclose(opt["y"][i], z[k]["shipped_opti"], rtol=0, atol=TOL)


def test_view_alignment_kernel_is_bit_identical_to_interp1d():
    """eksb_interp_rows == the oracle's paw_align (scipy interp1d arithmetic) bit for bit, incl. the x flip
    and ragged timestamp ranges."""
    from eks_b200.multiview_pca_smoother import _align_views
    from oracle import restated_smoothers as osm
    rng = np.random.default_rng(3)
    Tl, Tr, M = 400, 1000, 5
    ts_r = np.sort(rng.uniform(0.0, 10.0, Tr))
    ts_l = np.sort(rng.uniform(-0.5, 10.5, Tl))                      # some left frames fall outside the right range
    ts_l[50] = ts_r[123]                                             # an exact hit on a source timestamp
    left = rng.uniform(0, 128, size=(M, Tl, 6)); right = rng.uniform(0, 128, size=(M, Tr, 6))
    gl, gr = _align_views(left, right, ts_l, ts_r)
    for m in range(M):
        wl, wr = osm.paw_align(left[m], right[m], ts_l, ts_r)
        assert gl[m].shape == wl.shape and np.array_equal(gl[m], wl)
        assert np.array_equal(gr[m], wr)


def test_csv_to_csv_session_matches_shipped_eks_csv(tmp_path):
    """scripts/multicam_example.py end to end: member CSVs in -> eks.csv out, compared with the values the
    reference shipped in data/mirror-mouse/output/eks.csv (frozen in tests/golden/mirror_mouse.npz)."""
    from eks_b200 import io
    from eks_b200.session import smooth_multicam_csvs
    z = cases_of(load_golden("mirror_mouse.npz"))
    kps, cams = sorted(z), ["top", "bot"]
    M, T = z[kps[0]]["stack"].shape[:2]
    names = [f"{k}_{c}" for c in cams for k in kps]                  # file order: all top views, then all bot views
    cols = pd.MultiIndex.from_product([["tracker"], names, ["x", "y", "likelihood"]], names=["scorer", "bodyparts", "coords"])
    paths = []
    for m in range(M):
        body = np.full((T, len(cols)), 0.9)
        for k in kps:
            for ci, c in enumerate(cams):
                j = 3 * names.index(f"{k}_{c}")
                body[:, j:j + 2] = z[k]["stack"][m][:, 2 * ci:2 * ci + 2]
        path = tmp_path / f"member{m}.csv"
        pd.DataFrame(body, columns=cols).to_csv(path)
        paths.append(str(path))
    out = smooth_multicam_csvs(paths, kps, cams, 0.01, 25, save_path=str(tmp_path / "out" / "eks.csv"))
    back = pd.read_csv(tmp_path / "out" / "eks.csv", header=[0, 1, 2], index_col=0, float_precision="round_trip")
    assert back.columns[0][0] == "ensemble-kalman_tracker" and back.shape == (T, len(cols))
    for k in kps:
        got = np.column_stack([back[("ensemble-kalman_tracker", f"{k}_{c}", xy)].to_numpy() for c in cams for xy in ("x", "y")])
        np.testing.assert_allclose(got, z[k]["shipped_std"], rtol=0, atol=TOL)
        assert (back[("ensemble-kalman_tracker", f"{k}_top", "likelihood")] == 1.0).all()
