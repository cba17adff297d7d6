"""Developer probe: K8 (eksb_pca_setup) timing at bench scale on random obs records."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from eks_b200 import device as dev

def run(dt, B, T, p=4):
    d = torch.device("cuda", 0)
    g = torch.Generator(device=d); g.manual_seed(1)
    tiles = dev.num_tiles(B)
    obs = torch.empty(tiles, T, 2 * p, 32, dtype=dt, device=d)
    obs[:, :, :p] = torch.randn(tiles, T, p, 32, dtype=dt, device=d, generator=g).cumsum_(1)
    obs[:, :, p:] = torch.exp(torch.randn(tiles, T, p, 32, dtype=dt, device=d, generator=g) * 0.8 - 1.5)
    obs = obs.reshape(-1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for r in range(6):
        os.environ["EKSB_SETUP_NO_WIDE"] = "1" if r >= 3 else "0"      # last three: round-1 8-bit passes
        e0.record()
        prm, means, st = dev.pca_setup(obs, B, T, p, 25.0, 0.01, center=True)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(f"{str(dt):14s} B={B} T={T} {'narrow' if r >= 3 else 'wide  '} pca_setup {ms:8.2f} ms -> {B*T/ms*1e-6:7.2f} G tf/s, "
              f"{B*T*4*p*obs.element_size()/ms*1e-6:7.1f} GB/s algorithmic (4p per tf)", flush=True)
    assert int(st.sum()) == 0
    os.environ.pop("EKSB_SETUP_NO_WIDE", None)

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "long":      # few tracks, very long: under ncu for the per-kernel split
        run(torch.float64, 64, 2000000)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "bench":     # the bench's group shape, under ncu
        run(torch.float64, 8192, 100000)
        sys.exit(0)
    run(torch.float64, 8192, 20000)
    run(torch.float64, 8192, 100000)
    run(torch.float32, 16384, 20000)
    run(torch.float64, 64, 2000000)
