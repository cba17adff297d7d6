"""Developer probe: config-5 shape (few very long tracks) through the parallel-in-time scan."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eks_b200 import device as dev

def run(dt, B, T, chunk=0, reps=2):
    d = torch.device("cuda", 0)
    g = torch.Generator(device=d); g.manual_seed(1)
    p, n = 4, 3
    tiles = dev.num_tiles(B)
    obs = torch.empty(tiles, T, 2 * p, 32, dtype=dt, device=d)
    for t0 in range(0, T, 1_000_000):
        t1 = min(T, t0 + 1_000_000)
        obs[:, t0:t1, :p] = torch.randn(tiles, t1 - t0, p, 32, dtype=dt, device=d, generator=g)
        obs[:, t0:t1, p:] = torch.exp(torch.randn(tiles, t1 - t0, p, 32, dtype=dt, device=d, generator=g) * 0.8 - 1.5)
    obs = obs.reshape(-1)
    rng = np.random.default_rng(0)
    C = np.stack([np.linalg.qr(rng.normal(size=(p, n)))[0] for _ in range(B)])
    prm = dev.TrackParams.from_numpy(np.zeros(n), np.eye(n), C, np.eye(n), 1e-3 * np.eye(n), dtype=dt, device=d)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for r in range(reps + 1):
        e0.record()
        out = dev.kf_rts_scan(obs, prm, B, T, chunk=chunk)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if r:
            tf = B * T
            print(f"{str(dt):14s} B={B} T={T} chunk={out['chunk']} scan fwd+bwd {ms:8.2f} ms -> {tf/ms*1e-6:7.2f} G tf/s, "
                  f"{tf*(3*p+3*dev.sigma(n))*obs.element_size()/ms*1e-6:7.1f} GB/s algorithmic", flush=True)
        assert int(out["status"].sum()) == 0
        del out
    if T >= 10_000_000:
        for r in range(2):
            e0.record()
            out = dev.btd_solve(obs, prm, B, T)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            if r:
                print(f"{str(dt):14s} B={B} T={T} btd_solve (parallel-in-time route) {ms:8.2f} ms -> {B*T/ms*1e-6:7.2f} G tf/s", flush=True)
            del out
    torch.cuda.empty_cache()

if __name__ == "__main__":
    big_only = os.environ.get("SCAN_TIME_BIG_ONLY") == "1"
    dts = (torch.float64,) if os.environ.get("SCAN_TIME_F64_ONLY") == "1" else (torch.float32, torch.float64)
    for dt in dts:
        if not big_only:
            run(dt, 64, 2_000_000)
            run(dt, 64, 2_000_000, chunk=1024)
        run(dt, 64, 10_000_000)
