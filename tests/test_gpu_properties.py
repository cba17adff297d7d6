"""GPU: size-independent properties of the hot path at sizes the CPU oracle cannot reach
(thousands of tracks / a million frames), complementing the oracle comparisons in test_gpu_core.py:
  * the smoothed mean is LINEAR in the observations and the covariances do not depend on them
    (eks/ensemble_kalman.py:102-105,160-161: y only enters through the innovation);
  * the block-tridiagonal Newton solve equals filter + RTS;
  * the parallel-in-time scan equals the sequential kernels;
  * a shard of tracks processed alone equals the same tracks inside the full batch, bit for bit.
Tolerances: fp64 1e-8 (north_star bar 1e-5), fp32 1e-3 (the bar itself), each track's each component scaled by its own
largest magnitude over time (conftest.rel / rel_t), not by the largest number in the array."""
import numpy as np
import pytest

from conftest import rel, RTOL_F32

pytestmark = pytest.mark.gpu


def rel_t(got, want):
    """conftest.rel for track-tile tensors [tile][T][row][32]: per (tile, row, lane) scale = max over time."""
    got, want = got.double(), want.double()
    scale = want.abs().amax(dim=1, keepdim=True)
    floor = max(float(want.abs().max()) * 1e-6, 1e-300)
    return float(((got - want).abs() / scale.clamp_min(floor)).max())


def _batch(B, T, p=4, n=3, seed=0, dt=None):
    import torch
    from eks_b200 import device as dev
    dt = dt or torch.float64
    d = torch.device("cuda", 0)
    g = torch.Generator(device=d); g.manual_seed(seed)
    tiles = dev.num_tiles(B)
    obs = torch.empty(tiles, T, 2 * p, 32, dtype=dt, device=d)
    obs[:, :, :p] = torch.randn(tiles, T, p, 32, dtype=dt, device=d, generator=g).cumsum_(1).mul_(0.05)
    obs[:, :, p:] = torch.exp(torch.randn(tiles, T, p, 32, dtype=dt, device=d, generator=g) * 0.8 - 1.5)
    rng = np.random.default_rng(seed)
    C = np.linalg.qr(rng.normal(size=(B, p, n)))[0]
    prm = dev.TrackParams.from_numpy(np.zeros(n), np.eye(n), C, np.diag([1.0, 0.999, 0.99]), 1e-3 * np.eye(n), dtype=dt, device=d)
    return obs, prm


def _smooth(obs, prm, B, T):
    from eks_b200 import device as dev
    filt, _, st = dev.kf_forward(obs.reshape(-1), prm, B, T)
    out = dev.rts_backward(filt, prm, B, T, inplace=True, status=st)
    assert int(st.sum()) == 0
    return out["smth"].reshape(-1, T, 9, 32), out["ym"].reshape(-1, T, 4, 32)


def test_smoothed_mean_is_linear_in_y_and_covariance_ignores_y():
    import torch
    B, T = 4096, 5000
    obs1, prm = _batch(B, T, seed=1)
    obs2, _ = _batch(B, T, seed=2)
    obs2[:, :, 4:] = obs1[:, :, 4:]                                   # same ensemble variances
    a, b = 0.7, -1.9
    mix = obs1.clone(); mix[:, :, :4] = a * obs1[:, :, :4] + b * obs2[:, :, :4]
    s1, y1 = _smooth(obs1, prm, B, T); s2, y2 = _smooth(obs2, prm, B, T); sm, ym = _smooth(mix, prm, B, T)
    assert rel_t(a * s1[:, :, :3] + b * s2[:, :, :3], sm[:, :, :3]) < 1e-8
    assert rel_t(a * y1 + b * y2, ym) < 1e-8
    assert torch.equal(sm[:, :, 3:], s1[:, :, 3:]) and torch.equal(s2[:, :, 3:], s1[:, :, 3:])   # covariances: bit-identical


def test_block_tridiagonal_solve_equals_filter_plus_rts_at_scale():
    from eks_b200 import device as dev
    B, T = 2048, 20000
    obs, prm = _batch(B, T, seed=3)
    sm, _ = _smooth(obs, prm, B, T)
    sol = dev.btd_solve(obs.reshape(-1), prm, B, T, scan=False)               # the block-Thomas kernels
    assert int(sol["status"].sum()) == 0
    x = sol["sol"].reshape(-1, T, 9, 32)
    assert rel_t(x, sm) < 1e-8


def test_block_tridiagonal_solve_of_few_long_tracks_goes_parallel_in_time():
    """Config 5 shape: a handful of long tracks.  btd_solve routes them through the chunked scan; solution and
    diagonal blocks of H^-1 equal the block-Thomas kernels' (and, by the test above, filter + RTS)."""
    from eks_b200 import device as dev
    B, T = 64, 300_000
    obs, prm = _batch(B, T, seed=8)
    thomas = dev.btd_solve(obs.reshape(-1), prm, B, T, scan=False)
    auto = dev.btd_solve(obs.reshape(-1), prm, B, T)
    assert int(thomas["status"].sum()) == 0 and int(auto["status"].sum()) == 0
    a, b = auto["sol"].reshape(-1, T, 9, 32), thomas["sol"].reshape(-1, T, 9, 32)
    assert rel_t(a, b) < 1e-8


def test_scan_equals_sequential_on_long_tracks():
    from eks_b200 import device as dev
    B, T = 64, 1_000_000
    obs, prm = _batch(B, T, seed=4)
    sm, ym = _smooth(obs, prm, B, T)
    sc = dev.kf_rts_scan(obs.reshape(-1), prm, B, T)
    assert int(sc["status"].sum()) == 0
    assert rel_t(sc["smth"].reshape(-1, T, 9, 32), sm) < 1e-8
    assert rel_t(sc["ym"].reshape(-1, T, 4, 32), ym) < 1e-8


@pytest.mark.parametrize("early", ["1", "0"])
@pytest.mark.parametrize("dtname,tol", [("float64", 1e-8), ("float32", RTOL_F32)])
def test_scan_early_writeout_in_place_equals_sequential(dtname, tol, early, monkeypatch):
    """Chunks much longer than the filter's memory: the summary kernels write most records themselves
    (early write-out, smoothing in place over the filtered records) and the re-run kernels only the
    frames before each cut.  Same records as the sequential kernels, with the switch on and off."""
    import torch
    from eks_b200 import device as dev
    if early == "0":
        monkeypatch.setenv("EKSB_SCAN_NO_EARLY", "1")
    B, T = 96, 200_000
    obs, prm = _batch(B, T, seed=6, dt=getattr(torch, dtname))
    sm, ym = _smooth(obs, prm, B, T)
    sc = dev.kf_rts_scan(obs.reshape(-1), prm, B, T, chunk=10_000)
    assert int(sc["status"].sum()) == 0 and sc["smth"].data_ptr() == sc["filt"].data_ptr()
    assert rel_t(sc["smth"].reshape(-1, T, 9, 32), sm) < tol
    assert rel_t(sc["ym"].reshape(-1, T, 4, 32), ym) < tol


def test_scan_mixed_chunks_in_place_equals_sequential():
    """Some (tile, chunk) pairs never forget their start (a blind stretch of huge ensemble variance in one tile
    only), so the folds run and those chunks are re-run in full, while their neighbours take their start state
    from the records the summary kernels wrote.  Same records as the sequential kernels."""
    from eks_b200 import device as dev
    B, T, L = 96, 120_000, 10_000
    obs, prm = _batch(B, T, seed=7)
    obs[1, 39_000:71_000, 4:] *= 1e12                                  # tile 1: chunks 4..6 are blind
    sm, ym = _smooth(obs, prm, B, T)
    sc = dev.kf_rts_scan(obs.reshape(-1), prm, B, T, chunk=L)
    assert int(sc["status"].sum()) == 0
    assert rel_t(sc["smth"].reshape(-1, T, 9, 32), sm) < 1e-8          # per track: the blind tile has a far larger scale
    assert rel_t(sc["ym"].reshape(-1, T, 4, 32), ym) < 1e-8


def test_shard_of_a_large_batch_is_bitwise_identical():
    import torch
    from eks_b200 import device as dev
    B, T = 8192, 2000
    obs, prm = _batch(B, T, seed=5)
    full, _ = _smooth(obs, prm, B, T)
    lo, hi = 3072, 5120                                                # tile-aligned shard, as selection.shard_bounds yields
    sub = dev.TrackParams(prm.m0, prm.S0, prm.C[lo:hi].contiguous(), prm.A, prm.Q, a_mode=prm.a_mode)
    part, _ = _smooth(obs[lo // 32:hi // 32].contiguous(), sub, hi - lo, T)
    assert torch.equal(part, full[lo // 32:hi // 32])


@pytest.mark.parametrize("B", [148 * 32 + 7, 148 * 64 + 1, 148 * 96 - 31, 16384])
@pytest.mark.parametrize("dtname,tol", [("float64", 1e-9), ("float32", RTOL_F32)])
def test_every_tiles_per_cta_choice_matches_the_oracle(B, dtname, tol):
    """The launch packs ceil(tiles/148) tiles into a CTA (1..4) and places warp roles per scheduler; every
    choice must give the oracle's numbers (checked on a few tracks spread over the batch, ragged last tile)."""
    import torch
    from eks_b200 import device as dev
    from oracle import restated as o
    dt = getattr(torch, dtname)
    T, p, n = 61, 4, 3
    d = torch.device("cuda", 0)
    rng = np.random.default_rng(B)
    pick = sorted({0, 31, 32, B // 3, B // 2, B - 33, B - 1})
    C = np.linalg.qr(rng.normal(size=(B, p, n)))[0]
    A = np.diag([1.0, 0.99, 0.95]); Q = 0.01 * np.eye(n) + 0.001; S0 = np.diag([1.0, 2.0, 0.5]); m0 = np.array([0.3, -0.2, 0.1])
    y = rng.normal(size=(B, T, p)).cumsum(1) * 0.1
    ev = np.exp(rng.normal(np.log(0.05), 1.0, size=(B, T, p)))
    prm = dev.TrackParams.from_numpy(m0, S0, C, A, Q, dtype=dt, device=d)
    obs = dev.pack_obs(torch.as_tensor(y, dtype=dt).to(d), torch.as_tensor(ev, dtype=dt).to(d))
    filt, _, st = dev.kf_forward(obs, prm, B, T)
    out = dev.rts_backward(filt, prm, B, T, inplace=True, status=st)
    assert int(st.sum()) == 0
    sm = dev.unpack_rows(out["smth"], 9, 0, B, T, 9)[pick].cpu().numpy().astype(np.float64)
    ym = dev.unpack_rows(out["ym"], p, 0, B, T, p)[pick].cpu().numpy().astype(np.float64)
    iu = np.triu_indices(n)
    for i, b in enumerate(pick):
        yb, evb = (y[b], ev[b]) if dtname == "float64" else (y[b].astype(np.float32).astype(np.float64), ev[b].astype(np.float32).astype(np.float64))
        mf, Vf, S = o.filtering_pass(yb, m0, S0, C[b], np.eye(p), A, Q, evb)
        ms, Vs, _ = o.smooth_backward(yb, mf, Vf, S, A, Q, C[b])
        assert rel(sm[i][:, :n], ms) < tol and rel(sm[i][:, n:], Vs[:, iu[0], iu[1]]) < tol
        assert rel(ym[i], ms @ C[b].T) < tol


def test_nccl_grid_search_over_two_ranks():
    """The path's one collective on real GPUs: sharded NLL grid sums all-reduced over NCCL equal the
    single-rank sums (tools/nccl_selection_check.py).  Needs 2 GPUs; skipped on a 1-GPU box (the host-side
    logic is covered by the gloo world-2 test on CPU)."""
    import os, subprocess, sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29547", os.path.join(root, "tools", "nccl_selection_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and " OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
