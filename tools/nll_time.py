"""Developer probe: NLL grid sweep (A8) and z-score timing on the bench shape (device resident)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eks_b200 import device as dev, selection as sel

def run(dt, B, T, G, p=4, n=3):
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
    grid = np.logspace(-2, 2, G)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tots = {}
    for mode in ("1", "0"):                                  # pipelined NLL kernel / plain thread-per-track kernel
        os.environ["EKSB_NLL_PIPE"] = mode
        for r in range(3):
            e0.record()
            nll, st = sel.nll_grid(obs, prm, B, T, grid)
            tot = sel.nll_sums(nll)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            if r:
                print(f"{str(dt):14s} B={B} T={T} grid={G} {'pipe ' if mode == '1' else 'plain'}: NLL sweep {ms:8.2f} ms -> {B*T*G/ms*1e-6:7.2f} G track-frame-evaluations/s, "
                      f"obs bytes read once = {B*T*2*p*obs.element_size()/ms*1e-6:7.1f} GB/s", flush=True)
        tots[mode] = tot.cpu().numpy()
    os.environ.pop("EKSB_NLL_PIPE", None)
    print("   pipe vs plain: max rel diff of the grid sums", float(np.abs(tots["1"] - tots["0"]).max() / np.abs(tots["0"]).max()), flush=True)
    e0.record(); filt, _, _ = dev.kf_forward(obs, prm, B, T); e1.record(); torch.cuda.synchronize()
    e0.record(); filt, _, _ = dev.kf_forward(obs, prm, B, T, filt=filt); e1.record(); torch.cuda.synchronize()
    print(f"{str(dt):14s} B={B} T={T} forward filter (records written) {e0.elapsed_time(e1):8.2f} ms", flush=True)
    out = dev.rts_backward(filt, prm, B, T, inplace=True)
    for r in range(3):
        e0.record()
        z = sel.zscore(obs, out["smth"], prm, B, T)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if r:
            es = obs.element_size()
            print(f"{str(dt):14s} B={B} T={T} z-score {ms:8.2f} ms -> {B*T/ms*1e-6:7.2f} G tf/s, {B*T*(2*p+dev.sigma(n)+p)*es/ms*1e-6:7.1f} GB/s", flush=True)

if __name__ == "__main__":
    for dt in (torch.float64, torch.float32):
        run(dt, 8192, 20000, 16)
        run(dt, 8192, 20000, 1)
