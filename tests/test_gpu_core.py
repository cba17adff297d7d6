"""GPU parity tests: the CUDA path (through the C ABI) against golden fixtures and the oracle."""
import numpy as np
import pandas as pd
import pytest

from conftest import load_golden, cases_of, rel, RTOL_F32
from oracle import restated as o

pytestmark = pytest.mark.gpu

CORE = cases_of(load_golden("core_kats.npz"))
KF_CASES = [k for k in CORE if not k.startswith("ens_") and k != "denseAR"]
RTOL64 = 1e-5   # north_star: smoothed means/variances within 1e-5 relative in fp64 (we get ~1e-13)


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


def _random_batch(B, T, p, n, dt, seed=0, a_diag=True):
    import torch
    from eks_b200 import device as dev
    rng = np.random.default_rng(seed)
    C = np.stack([np.linalg.qr(rng.normal(size=(p, n)))[0] for _ in range(B)])
    A = np.diag(rng.uniform(0.95, 1.0, n)) if a_diag else np.eye(n) * 0.97 + 0.01 * rng.normal(size=(n, n))
    Q = 0.01 * np.eye(n) + 0.001
    S0 = np.diag(rng.uniform(0.5, 2, n)); m0 = rng.normal(size=n)
    x = np.cumsum(rng.normal(size=(B, T, n)) * 0.1, axis=1)
    sd = np.exp(rng.normal(np.log(0.5), 0.8, size=(B, T, p)))
    y = np.einsum("bpn,btn->btp", C, x) + sd * rng.normal(size=(B, T, p))
    ev = sd ** 2 / 5
    d = torch.device("cuda", 0)
    prm = dev.TrackParams.from_numpy(m0, S0, C, A, Q, dtype=dt, device=d)
    obs = dev.pack_obs(torch.as_tensor(y, dtype=dt).to(d), torch.as_tensor(ev, dtype=dt).to(d))
    return dict(y=y, ev=ev, C=C, A=A, Q=Q, S0=S0, m0=m0, prm=prm, obs=obs)


@pytest.mark.parametrize("n,p", [(3, 4), (3, 6), (3, 8), (2, 4), (4, 6), (6, 8)])
@pytest.mark.parametrize("dtname,tol", [("float64", 1e-9), ("float32", 1e-3)])
@pytest.mark.parametrize("a_diag", [True, False])
def test_batched_pipe_and_plain_kernels_vs_oracle(n, p, dtname, tol, a_diag, monkeypatch):
    """Ragged batch (B not a multiple of 32, T not a multiple of the hand-off batch): both the
    warp-specialised and the plain kernels against the oracle, per track."""
    import torch
    from eks_b200 import device as dev
    dt = getattr(torch, dtname)
    B, T = 70, 203
    c = _random_batch(B, T, p, n, dt, seed=n * 10 + p, a_diag=a_diag)
    res = {}
    for label, want_nll in (("pipe", False), ("plain", True)):     # an NLL request routes to the plain kernel
        filt, nll, st = dev.kf_forward(c["obs"], c["prm"], B, T, want_nll=want_nll)
        out = dev.rts_backward(filt, c["prm"], B, T, want_ym=True, want_yv=True, want_cv=want_nll)
        assert int(st.sum()) == 0 and int(out["status"].sum()) == 0
        sg = dev.sigma(n)
        res[label] = dict(
            mf=dev.unpack_rows(filt, sg, 0, B, T, n).cpu().numpy().astype(np.float64),
            ms=dev.unpack_rows(out["smth"], sg, 0, B, T, n).cpu().numpy().astype(np.float64),
            Vsp=dev.unpack_rows(out["smth"], sg, n, B, T, sg - n).cpu().numpy().astype(np.float64),
            ym=dev.unpack_rows(out["ym"], p, 0, B, T, p).cpu().numpy().astype(np.float64),
            yv=dev.unpack_rows(out["yv"], p, 0, B, T, p).cpu().numpy().astype(np.float64))
    iu = np.triu_indices(n)
    for b in (0, 31, 32, B - 1):
        yq = c["y"].astype(np.float32).astype(np.float64) if dtname == "float32" else c["y"]
        evq = c["ev"].astype(np.float32).astype(np.float64) if dtname == "float32" else c["ev"]
        mf, Vf, S = o.filtering_pass(yq[b], c["m0"], c["S0"], c["C"][b], np.eye(p), c["A"], c["Q"], evq[b])
        ms, Vs, _ = o.smooth_backward(yq[b], mf, Vf, S, c["A"], c["Q"], c["C"][b])
        for label in ("pipe", "plain"):
            r = res[label]
            assert rel(r["mf"][b], mf) < tol, (label, b)
            assert rel(r["ms"][b], ms) < tol, (label, b)
            assert rel(r["Vsp"][b], Vs[:, iu[0], iu[1]]) < tol, (label, b)
            assert rel(r["ym"][b], ms @ c["C"][b].T) < tol, (label, b)
            assert rel(r["yv"][b], np.einsum("jk,tkl,jl->tj", c["C"][b], Vs, c["C"][b])) < tol, (label, b)


@pytest.mark.parametrize("dtname,tol", [("float64", 1e-11), ("float32", 1e-4)])
def test_large_batches_go_out_as_groups_of_tiles(dtname, tol, monkeypatch):
    """A batch with more tiles than the pipelined kernels handle at their best is launched as several
    groups of tiles (eksb_kf.cuh launch_forward / launch_backward).  Forced here on 7 ragged tiles with
    per-track C: same records as one launch and as the plain kernels, status lands on the right track."""
    import torch
    from eks_b200 import device as dev
    dt = getattr(torch, dtname)
    B, T, p, n = 200, 150, 4, 3
    c = _random_batch(B, T, p, n, dt, seed=77)

    def run():
        filt, _, st = dev.kf_forward(c["obs"], c["prm"], B, T)
        out = dev.rts_backward(filt, c["prm"], B, T, want_ym=True, want_yv=True)
        btd = dev.btd_solve(c["obs"], c["prm"], B, T, want_ym=True, scan=False)
        assert int(btd["status"].sum()) == 0
        return filt.clone(), out["smth"], out["ym"], out["yv"], st, out["status"], btd["sol"], btd["ym"]

    one = run()
    monkeypatch.setenv("EKSB_PIPE_MAX_TILES", "2")
    monkeypatch.setenv("EKSB_PIPE_GROUP_TILES", "3")                   # 7 tiles -> groups of 3, 3, 1
    grouped = run()
    monkeypatch.setenv("EKSB_PIPE", "0")
    plain = run()
    tol_plain = 1e-9 if dtname == "float64" else RTOL_F32                  # different arithmetic order (north_star bars 1e-5 / 1e-3)
    def rel_t(x, ref):                                                 # conftest.rel on [tile][T][row][32] records
        x, ref = x.double().reshape(-1, T, x.numel() // (dev.num_tiles(B) * T * 32), 32), ref.double().reshape(-1, T, ref.numel() // (dev.num_tiles(B) * T * 32), 32)
        scale = ref.abs().amax(dim=1, keepdim=True).clamp_min(float(ref.abs().max()) * 1e-6)
        return float(((x - ref).abs() / scale).max())
    for a, b, e in zip(one[:4], grouped[:4], plain[:4]):
        assert rel_t(b, a) <= tol
        assert rel_t(e, b) <= tol_plain
    for a, b in zip(one[6:], grouped[6:]):                             # block-tridiagonal solve, same grouping
        assert rel_t(b, a) <= tol
    assert rel_t(grouped[6], grouped[1]) <= tol_plain                  # == filter + RTS
    assert int(grouped[4].sum()) == 0 and int(grouped[5].sum()) == 0
    # a track with a non-positive ensemble variance in the last group reports through its own status slot
    monkeypatch.setenv("EKSB_PIPE", "1")
    bad = 197
    obs = c["obs"].clone().reshape(-1, T, 2 * p, 32)
    obs[bad // 32, 10, p:, bad % 32] = -1e-6           # det(Id + Lam S) < 0 at that frame
    _, _, st = dev.kf_forward(obs.reshape(-1), c["prm"], B, T)
    assert int(st[bad]) & 2 and int((st != 0).sum()) == 1              # EKSB_STATUS_NONPOS_VAR, in the last group's slot


@pytest.mark.parametrize("dtname", ["float64", "float32"])
@pytest.mark.parametrize("B", [70, 200])
def test_observation_offset_equals_centring_pass(dtname, B, monkeypatch):
    """eksb_kf_forward_offset: y - offset formed as the frames are read == filtering records centred by a separate
    pass, bit for bit, in the pipelined kernels (three-role and two-role, grouped launches) and the plain kernel
    (with NLL)."""
    import torch
    from eks_b200 import device as dev
    dt = getattr(torch, dtname)
    T, p, n = 260, 4, 3
    c = _random_batch(B, T, p, n, dt, seed=5 + B)
    d = c["obs"].device
    off = torch.as_tensor(np.random.default_rng(B).uniform(-300, 300, size=(B, p)), dtype=dt).to(d)
    raw = dev.pack_obs(torch.as_tensor(c["y"], dtype=dt).to(d) + off[:, None, :], torch.as_tensor(c["ev"], dtype=dt).to(d))
    y_c = (torch.as_tensor(c["y"], dtype=dt).to(d) + off[:, None, :]) - off[:, None, :]       # the centring pass's rounding
    centred = dev.pack_obs(y_c, torch.as_tensor(c["ev"], dtype=dt).to(d))
    rows = lambda rec: dev.unpack_rows(rec, dev.sigma(n), 0, B, T, dev.sigma(n))             # real tracks only (not the padding lanes)
    for env in ({}, {"EKSB_PIPE_MAX_TILES": "2", "EKSB_PIPE_GROUP_TILES": "3"}, {"EKSB_PIPE": "0"}):
        for k in ("EKSB_PIPE_MAX_TILES", "EKSB_PIPE_GROUP_TILES", "EKSB_PIPE"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        a, _, sa = dev.kf_forward(centred, c["prm"], B, T)
        b, _, sb = dev.kf_forward(raw, c["prm"], B, T, y_offset=off)
        assert torch.equal(rows(a), rows(b)) and int(sa.sum()) == 0 and int(sb.sum()) == 0
    a, na, _ = dev.kf_forward(centred, c["prm"], B, T, want_nll=True)                        # plain kernel, NLL
    b, nb, _ = dev.kf_forward(raw, c["prm"], B, T, want_nll=True, y_offset=off)
    assert torch.equal(rows(a), rows(b)) and torch.equal(na, nb)


def test_inplace_smoothing_equals_out_of_place():
    import torch
    from eks_b200 import device as dev
    c = _random_batch(40, 150, 4, 3, torch.float64, seed=3)
    filt, _, _ = dev.kf_forward(c["obs"], c["prm"], 40, 150)
    a = dev.rts_backward(filt, c["prm"], 40, 150)
    b = dev.rts_backward(filt.clone(), c["prm"], 40, 150, inplace=True)
    assert torch.equal(a["smth"], b["smth"]) and torch.equal(a["ym"], b["ym"])


@pytest.mark.parametrize("name", ["p4", "p6", "p8", "n2p4", "n4p6", "T2", "T1"])
def test_block_tridiagonal_solve_equals_rts(name):
    """A7 through the reference-named entry point: Newton step == RTS mean, H^-1 blocks == Vs."""
    from eks_b200.block_tridiag import Block_thomas_algo
    c = CORE[name]
    s, cov = Block_thomas_algo(c["y"], c["m0"], c["S0"], c["A"], c["C"], c["Q"], c["ev"], return_cov=True)
    want_m, want_V = (c["ms"], c["Vs"]) if "ms" in c else (c["mf"], c["Vf"])
    assert rel(s, want_m) < 1e-9 and rel(cov, want_V) < 1e-9
    assert rel(s, o.btd_solve(c["y"], c["m0"], c["S0"], c["A"], c["C"], c["Q"], c["ev"])) < 1e-9


def test_block_tridiagonal_batched_f32_and_dense_A():
    import torch
    from eks_b200 import device as dev
    for dtname, tol, a_diag in (("float32", RTOL_F32, True), ("float64", 1e-9, False)):
        dt = getattr(torch, dtname)
        B, T, p, n = 45, 131, 4, 3
        c = _random_batch(B, T, p, n, dt, seed=11, a_diag=a_diag)
        out = dev.btd_solve(c["obs"], c["prm"], B, T, want_ym=True)
        assert int(out["status"].sum()) == 0
        x = dev.unpack_rows(out["sol"], dev.sigma(n), 0, B, T, n).cpu().numpy().astype(np.float64)
        ymv = dev.unpack_rows(out["ym"], p, 0, B, T, p).cpu().numpy().astype(np.float64)
        for b in (0, 33, B - 1):
            yq = c["y"].astype(np.float32).astype(np.float64) if dtname == "float32" else c["y"]
            evq = c["ev"].astype(np.float32).astype(np.float64) if dtname == "float32" else c["ev"]
            want = o.btd_solve(yq[b], c["m0"], c["S0"], c["A"], c["C"][b], c["Q"], evq[b])
            assert rel(x[b], want) < tol and rel(ymv[b], want @ c["C"][b].T) < tol


def test_nll_grid_zscore_and_selection():
    import torch
    from eks_b200 import device as dev, selection as sel
    B, T, p, n = 37, 120, 4, 3
    c = _random_batch(B, T, p, n, torch.float64, seed=21)
    grid = [0.1, 0.5, 1.0, 4.0, 20.0]
    nll, st = sel.nll_grid(c["obs"], c["prm"], B, T, grid)
    assert nll.shape == (len(grid), B) and int(st.sum()) == 0
    nll_h = nll.cpu().numpy()
    for gi, s in enumerate(grid):
        for b in (0, 36):
            want = o.nll(c["y"][b], c["m0"], c["S0"], c["C"][b], np.eye(p), c["A"], s * c["Q"], c["ev"][b])
            assert abs(nll_h[gi, b] - want) < 1e-8 * abs(want)
    sums = sel.nll_sums(nll).cpu().numpy()
    np.testing.assert_allclose(sums, nll_h.sum(axis=1), rtol=1e-12)
    best, tot = sel.select_smooth_param(c["obs"], c["prm"], B, T, grid)
    assert best == grid[int(np.argmin(nll_h.sum(axis=1)))]
    # z-score of the smoothed path
    filt, _, _ = dev.kf_forward(c["obs"], c["prm"], B, T)
    out = dev.rts_backward(filt, c["prm"], B, T)
    for add_ev in (True, False):
        z = dev.unpack_rows(sel.zscore(c["obs"], out["smth"], c["prm"], B, T, add_ev), p, 0, B, T, p).cpu().numpy()
        b = 5
        mf, Vf, S = o.filtering_pass(c["y"][b], c["m0"], c["S0"], c["C"][b], np.eye(p), c["A"], c["Q"], c["ev"][b])
        ms, Vs, _ = o.smooth_backward(c["y"][b], mf, Vf, S, c["A"], c["Q"], c["C"][b])
        np.testing.assert_allclose(z[b], o.zscore(c["y"][b], ms, Vs, c["C"][b], c["ev"][b], add_ev), rtol=1e-8, atol=1e-10)


def test_track_sharding_is_bitwise_identical_to_one_batch():
    """(e) multi-GPU: tracks are independent, so a shard processed alone gives bit-identical results
    to the same tracks inside the full batch, and shard NLL sums add up to the global sum."""
    import torch
    from eks_b200 import device as dev, selection as sel
    B, T, p, n = 96, 90, 4, 3
    c = _random_batch(B, T, p, n, torch.float64, seed=31)
    full, _, _ = dev.kf_forward(c["obs"], c["prm"], B, T)
    full = dev.rts_backward(full, c["prm"], B, T, inplace=True)["smth"]
    ms_full = dev.unpack_rows(full, dev.sigma(n), 0, B, T, dev.sigma(n))
    grid = [0.3, 1.0, 3.0]
    tot_full = sel.nll_sums(sel.nll_grid(c["obs"], c["prm"], B, T, grid)[0]).cpu().numpy()
    acc = np.zeros(len(grid))
    for rank in range(3):
        lo, hi = sel.shard_bounds(B, rank, 3)
        assert lo % 32 == 0
        d = c["obs"].device
        prm = dev.TrackParams.from_numpy(c["m0"], c["S0"], c["C"][lo:hi], c["A"], c["Q"], device=d)
        obs = dev.pack_obs(torch.as_tensor(c["y"][lo:hi]).to(d), torch.as_tensor(c["ev"][lo:hi]).to(d))
        f, _, _ = dev.kf_forward(obs, prm, hi - lo, T)
        s = dev.rts_backward(f, prm, hi - lo, T, inplace=True)["smth"]
        got = dev.unpack_rows(s, dev.sigma(n), 0, hi - lo, T, dev.sigma(n))
        assert torch.equal(got, ms_full[lo:hi])
        acc += sel.nll_sums(sel.nll_grid(obs, prm, hi - lo, T, grid)[0]).cpu().numpy()
    np.testing.assert_allclose(acc, tot_full, rtol=1e-12)


def test_batched_host_api_matches_oracle():
    import torch
    from eks_b200 import batched
    B, T, p, n = 70, 140, 4, 3
    c = _random_batch(B, T, p, n, torch.float64, seed=41)
    cnt = {}
    out = batched.smooth_tracks(c["y"], c["ev"], c["m0"], c["S0"], c["C"], c["A"], c["Q"], block_tracks=32, counters=cnt,
                                outputs=("ms", "Vs", "ym"))
    assert cnt["launches"] == 9 * 3 and cnt["h2d_bytes"] == 2 * B * T * p * 8
    assert cnt["d2h_bytes"] == B * T * (3 + 6 + 4) * 8 + B * 4
    iu = np.triu_indices(n)
    for b in (0, 31, 32, 69):
        mf, Vf, S = o.filtering_pass(c["y"][b], c["m0"], c["S0"], c["C"][b], np.eye(p), c["A"], c["Q"], c["ev"][b])
        ms, Vs, _ = o.smooth_backward(c["y"][b], mf, Vf, S, c["A"], c["Q"], c["C"][b])
        assert rel(out["ms"][b].numpy(), ms) < 1e-9 and rel(out["Vs"][b].numpy(), Vs[:, iu[0], iu[1]]) < 1e-9
        assert rel(out["ym"][b].numpy(), ms @ c["C"][b].T) < 1e-9
    assert int(out["status"].sum()) == 0


def test_batched_host_api_scan_path_matches_oracle():
    """smooth_tracks routes small blocks of long tracks through the parallel-in-time scan (the e2e
    bench path); forced here on a short case and compared with the oracle and the sequential path."""
    import torch
    from eks_b200 import batched
    B, T, p, n = 40, 4500, 4, 3
    c = _random_batch(B, T, p, n, torch.float64, seed=43)
    full = ("ms", "Vs", "ym")
    auto = batched.smooth_tracks(c["y"], c["ev"], c["m0"], c["S0"], c["C"], c["A"], c["Q"], block_tracks=32, outputs=full)   # T >= 4096 -> scan
    seq = batched.smooth_tracks(c["y"], c["ev"], c["m0"], c["S0"], c["C"], c["A"], c["Q"], block_tracks=32, scan=False, outputs=full)
    iu = np.triu_indices(n)
    for k in full:
        assert rel(auto[k].numpy(), seq[k].numpy(), time_axis=1) < 1e-9
    for b in (0, 39):
        mf, Vf, S = o.filtering_pass(c["y"][b], c["m0"], c["S0"], c["C"][b], np.eye(p), c["A"], c["Q"], c["ev"][b])
        ms, Vs, _ = o.smooth_backward(c["y"][b], mf, Vf, S, c["A"], c["Q"], c["C"][b])
        assert rel(auto["ms"][b].numpy(), ms) < 1e-9 and rel(auto["Vs"][b].numpy(), Vs[:, iu[0], iu[1]]) < 1e-9
    assert int(auto["status"].sum()) == 0


@pytest.mark.parametrize("dtname,tol", [("float64", 1e-9), ("float32", RTOL_F32)])
@pytest.mark.parametrize("B,T,chunk", [(3, 1000, 64), (40, 777, 100), (2, 5000, 0), (33, 129, 200)])
def test_parallel_in_time_scan_equals_sequential(dtname, tol, B, T, chunk):
    """K4: the chunked scan reproduces the sequential kernels (and through them the oracle),
    for ragged last chunks, auto chunk length and a single chunk (chunk > T)."""
    import torch
    from eks_b200 import device as dev
    dt = getattr(torch, dtname)
    p, n = 4, 3
    c = _random_batch(B, T, p, n, dt, seed=B + T)
    filt, _, _ = dev.kf_forward(c["obs"], c["prm"], B, T)
    mf_seq = dev.unpack_rows(filt, dev.sigma(n), 0, B, T, dev.sigma(n)).cpu().numpy().astype(np.float64)
    seq = dev.rts_backward(filt, c["prm"], B, T)
    ms_seq = dev.unpack_rows(seq["smth"], dev.sigma(n), 0, B, T, dev.sigma(n)).cpu().numpy().astype(np.float64)
    ym_seq = dev.unpack_rows(seq["ym"], p, 0, B, T, p).cpu().numpy().astype(np.float64)
    f_only = dev.kf_rts_scan(c["obs"], c["prm"], B, T, chunk=chunk, smooth=False)
    mf_scan = dev.unpack_rows(f_only["filt"], dev.sigma(n), 0, B, T, dev.sigma(n)).cpu().numpy().astype(np.float64)
    assert rel(mf_scan, mf_seq, time_axis=1) < tol
    out = dev.kf_rts_scan(c["obs"], c["prm"], B, T, chunk=chunk)
    assert int(out["status"].sum()) == 0
    ms_scan = dev.unpack_rows(out["smth"], dev.sigma(n), 0, B, T, dev.sigma(n)).cpu().numpy().astype(np.float64)
    ym_scan = dev.unpack_rows(out["ym"], p, 0, B, T, p).cpu().numpy().astype(np.float64)
    assert rel(ms_scan, ms_seq, time_axis=1) < tol and rel(ym_scan, ym_seq, time_axis=1) < tol
    if dtname == "float64":
        b = B - 1
        mf, Vf, S = o.filtering_pass(c["y"][b], c["m0"], c["S0"], c["C"][b], np.eye(p), c["A"], c["Q"], c["ev"][b])
        ms, Vs, _ = o.smooth_backward(c["y"][b], mf, Vf, S, c["A"], c["Q"], c["C"][b])
        assert rel(ms_scan[b][:, :n], ms) < 1e-8


@pytest.mark.parametrize("pipe", ["1", "0"])
def test_nan_propagates_like_the_reference_and_stays_in_its_track(pipe, monkeypatch):
    """A NaN observation poisons the means of ITS track from that frame on and leaves the covariances alone
    (y only enters the innovation, eks/ensemble_kalman.py:102-105); a NaN ensemble variance poisons both.
    Neither raises in the reference (numpy.linalg.solve returns NaN), so neither sets `status`; other tracks
    of the tile are untouched.  Same NaN pattern as the oracle, frame by frame."""
    import torch
    from eks_b200 import device as dev
    monkeypatch.setenv("EKSB_PIPE", pipe)
    B, T, p, n = 40, 90, 4, 3
    c = _random_batch(B, T, p, n, torch.float64, seed=77)
    y, ev = c["y"].copy(), c["ev"].copy()
    y[3, 20, 1] = np.nan
    ev[35, 50, 2] = np.nan
    d = torch.device("cuda", 0)
    obs = dev.pack_obs(torch.as_tensor(y).to(d), torch.as_tensor(ev).to(d))
    filt, _, st = dev.kf_forward(obs, c["prm"], B, T)
    out = dev.rts_backward(filt, c["prm"], B, T, inplace=False, status=st)
    assert int(st.sum()) == 0
    mfV = dev.unpack_rows(filt, 9, 0, B, T, 9).cpu().numpy()
    sm = dev.unpack_rows(out["smth"], 9, 0, B, T, 9).cpu().numpy()
    iu = np.triu_indices(n)
    for b in (3, 35, 4, 34):
        mf, Vf, S = o.filtering_pass(y[b], c["m0"], c["S0"], c["C"][b], np.eye(p), c["A"], c["Q"], ev[b])
        ms, Vs, _ = o.smooth_backward(y[b], mf, Vf, S, c["A"], c["Q"], c["C"][b])
        for got, want in ((mfV[b][:, :n], mf), (mfV[b][:, n:], Vf[:, iu[0], iu[1]]), (sm[b][:, :n], ms), (sm[b][:, n:], Vs[:, iu[0], iu[1]])):
            assert np.array_equal(np.isnan(got), np.isnan(want))
            np.testing.assert_allclose(got[~np.isnan(want)], want[~np.isnan(want)], rtol=1e-9, atol=1e-12)
    assert np.isnan(mfV[3][20:, :n]).all() and not np.isnan(mfV[3][:, n:]).any() and not np.isnan(mfV[4]).any()
