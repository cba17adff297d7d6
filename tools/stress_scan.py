"""Randomised sweep (developer probe) of the parallel-in-time scan against the sequential kernels on the device:
random track counts, lengths, chunk lengths, process noise over six decades (so that chunks forget their start
quickly, slowly or never), blind stretches of huge ensemble variance in random tiles, both dtypes, in place."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from eks_b200 import device as dev

rng = np.random.default_rng(int(os.environ.get("SEED", "0")))
d = torch.device("cuda", 0)
worst = {torch.float64: 0.0, torch.float32: 0.0}
nfail = 0
iters = int(os.environ.get("ITERS", "120"))
for it in range(iters):
    n, p = [(3, 4), (3, 6), (3, 8), (2, 4)][rng.integers(4)]
    B = int(rng.choice([1, 7, 32, 33, 64, 97, 130]))
    T = int(rng.integers(2000, 60000))
    chunk = int(rng.choice([0, 300, 777, 1500, 4000, 9000]))
    dt, tol = (torch.float64, 1e-8) if rng.random() < 0.6 else (torch.float32, 5e-3)
    q = 10.0 ** rng.uniform(-7, -1)
    tiles = dev.num_tiles(B)
    g = torch.Generator(device=d); g.manual_seed(int(rng.integers(1 << 30)))
    obs = torch.empty(tiles, T, 2 * p, 32, dtype=dt, device=d)
    obs[:, :, :p] = torch.randn(tiles, T, p, 32, dtype=dt, device=d, generator=g).cumsum_(1).mul_(0.05).add_(float(rng.uniform(-300, 300)))
    obs[:, :, p:] = torch.exp(torch.randn(tiles, T, p, 32, dtype=dt, device=d, generator=g) * 0.8 - 1.5)
    if rng.random() < 0.4:                                   # a blind stretch in one tile
        t0 = int(rng.integers(0, T - 100)); t1 = min(T, t0 + int(rng.integers(100, 12000)))
        obs[int(rng.integers(tiles)), t0:t1, p:] *= (1e12 if dt == torch.float64 else 1e6)
    C = np.linalg.qr(rng.normal(size=(B, p, n)))[0]
    A = np.diag(rng.uniform(0.97, 1.0, n))
    prm = dev.TrackParams.from_numpy(rng.normal(size=n), np.eye(n), C, A, q * np.eye(n), dtype=dt, device=d)
    filt, _, st = dev.kf_forward(obs.reshape(-1), prm, B, T)
    seq = dev.rts_backward(filt, prm, B, T, inplace=True, status=st)
    sc = dev.kf_rts_scan(obs.reshape(-1), prm, B, T, chunk=chunk)
    sg = dev.sigma(n)
    a, b = sc["smth"].reshape(-1, T, sg, 32), seq["smth"].reshape(-1, T, sg, 32)
    err = 0.0
    for gi in range(tiles):                                  # per tile: a blind tile has a far larger covariance scale
        for lo, hi in ((0, n), (n, sg)):
            scale = float(b[gi, :, lo:hi].abs().max())
            err = max(err, float((a[gi, :, lo:hi] - b[gi, :, lo:hi]).abs().max()) / scale)
    ya, yb = sc["ym"], seq["ym"]
    err = max(err, float((ya - yb).abs().max()) / float(yb.abs().max()))
    bad = not (err < tol) or int(sc["status"].sum()) != int(st.sum())
    worst[dt] = max(worst[dt], err)
    if bad:
        nfail += 1
        print(f"FAIL it={it} n={n} p={p} B={B} T={T} chunk={chunk} q={q:.2e} {dt} err={err:.3e}", flush=True)
    del obs, filt, seq, sc
print(f"{iters} cases, {nfail} failures; worst relative difference f64 {worst[torch.float64]:.2e}, f32 {worst[torch.float32]:.2e}")
sys.exit(1 if nfail else 0)
