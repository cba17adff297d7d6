"""Developer probe: NLL sweep, plain launch vs time-split (eksb_kf_nll_ws), few tiles / long tracks and the
single-evaluation case at bench scale."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eks_b200 import device as dev


def run(dt, B, T, G, p=4, n=3):
    d = torch.device("cuda", 0)
    g = torch.Generator(device=d); g.manual_seed(1)
    tiles = dev.num_tiles(B)
    obs = torch.empty(tiles, T, 2 * p, 32, dtype=dt, device=d)
    obs[:, :, :p] = torch.randn(tiles, T, p, 32, dtype=dt, device=d, generator=g)
    obs[:, :, p:] = torch.exp(torch.randn(tiles, T, p, 32, dtype=dt, device=d, generator=g) * 0.8 - 1.5)
    obs = obs.reshape(-1)
    rng = np.random.default_rng(0)
    C = np.linalg.qr(rng.normal(size=(B, p, n)))[0]
    prm = dev.TrackParams.from_numpy(np.zeros(n), np.eye(n), C, np.eye(n), 1e-3 * np.eye(n), dtype=dt, device=d)
    q = None if G == 0 else torch.as_tensor(np.logspace(-2, 1, G), dtype=dt).to(d)
    ws = dev.nll_workspace(B, T, n, p, G, dt, d)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    res = {}
    for name, w in (("plain", None), ("split", ws)):
        if name == "split" and ws is None:
            continue
        ms = []
        for r in range(4):
            e0.record()
            a, st = dev.kf_nll(obs, prm, B, T, q_scale=q, workspace=w)
            e1.record(); torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        res[name] = (min(ms[1:]), a)
    line = f"{str(dt):14s} B={B:5d} T={T:8d} scales={max(G,1):2d}  plain {res['plain'][0]:8.2f} ms"
    if "split" in res:
        rel = float(((res["split"][1] - res["plain"][1]).abs() / res["plain"][1].abs()).max())
        line += f"  time-split {res['split'][0]:8.2f} ms  ({res['plain'][0] / res['split'][0]:5.1f}x, max rel diff {rel:.1e}, workspace {ws.numel() / 1e6:.1f} MB)"
    else:
        line += "  (not routed)"
    print(line, flush=True)


if __name__ == "__main__":
    for dt in (torch.float64, torch.float32):
        run(dt, 1, 100000, 0)          # one keypoint under an optimiser
        run(dt, 1, 100000, 16)
        run(dt, 32, 100000, 16)
        run(dt, 64, 1000000, 16)
        run(dt, 512, 100000, 16)
        run(dt, 2048, 100000, 16)
        run(dt, 8192, 20000, 0)        # single evaluation at bench scale
        run(dt, 8192, 20000, 16)
