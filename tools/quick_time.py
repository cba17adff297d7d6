"""Developer probe: device-resident filter+RTS timing on random track-tile data (not the bench)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eks_b200 import device as dev

def run(dt, B, T, p=4, n=3, reps=3):
    d = torch.device("cuda", 0)
    g = torch.Generator(device=d); g.manual_seed(1)
    tiles = dev.num_tiles(B)
    obs = torch.empty(tiles, T, 2 * p, 32, dtype=dt, device=d)
    obs[:, :, :p] = torch.randn(tiles, T, p, 32, dtype=dt, device=d, generator=g)
    obs[:, :, p:] = torch.exp(torch.randn(tiles, T, p, 32, dtype=dt, device=d, generator=g) * 0.8 - 1.5)
    obs = obs.reshape(-1)
    rng = np.random.default_rng(0)
    C = np.stack([np.linalg.qr(rng.normal(size=(p, n)))[0] for _ in range(B)])
    prm = dev.TrackParams.from_numpy(np.zeros(n), np.eye(n), C, np.eye(n), 1e-3 * np.eye(n), dtype=dt, device=d)
    filt = dev.alloc_records(B, T, dev.sigma(n), dt, d)
    ym = dev.alloc_records(B, T, p, dt, d)
    es = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    for r in range(reps + 1):
        es[0].record()
        dev.kf_forward(obs, prm, B, T, filt=filt)
        es[1].record()
        dev.rts_backward(filt, prm, B, T, inplace=True, ym=ym)
        es[2].record()
        torch.cuda.synchronize()
        f, b = es[0].elapsed_time(es[1]), es[1].elapsed_time(es[2])
        if r:
            bytes_tf = (3 * p + 3 * dev.sigma(n)) * obs.element_size()
            tf = B * T
            print(f"{str(dt):14s} B={B} T={T} fwd {f:8.2f} ms  bwd {b:8.2f} ms  -> {tf/(f+b)*1e-6:8.2f} G tf/s, "
                  f"{tf*bytes_tf/(f+b)*1e-6:7.1f} GB/s algorithmic (fwd {tf*(2*p+dev.sigma(n))*obs.element_size()/f*1e-6:7.1f}, "
                  f"bwd {tf*(2*dev.sigma(n)+p)*obs.element_size()/b*1e-6:7.1f})", flush=True)
    assert torch.isfinite(ym).all()

if __name__ == "__main__":
    if len(sys.argv) > 1:                       # bits:B:T ...
        for a in sys.argv[1:]:
            bits, B, T = (int(x) for x in a.split(":"))
            run(torch.float64 if bits == 64 else torch.float32, B, T, reps=int(os.environ.get("REPS", "2")))
    else:
        for dt in (torch.float32, torch.float64):
            for B, T in ((16384, 10000), (8192, 10000), (49152, 4000)):
                run(dt, B, T)
