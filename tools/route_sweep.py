"""Developer probe: sequential pipelined kernels vs the parallel-in-time scan as a function of the
number of track tiles (the routing policy of eksb_kf_forward_ws / eksb_rts_backward_ws and the
strong-scaling shards of bench.py: 16 384 tracks over 1/2/4/8 GPUs = 512/256/128/64 tiles)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from eks_b200 import device as dev


def make_obs(dt, tiles, T, p, d):
    g = torch.Generator(device=d); g.manual_seed(1)
    obs = torch.empty(tiles, T, 2 * p, 32, dtype=dt, device=d)
    step = max(1, (1 << 28) // (T * p * 32))
    for t0 in range(0, tiles, step):
        t1 = min(tiles, t0 + step)
        obs[t0:t1, :, :p] = torch.randn(t1 - t0, T, p, 32, dtype=dt, device=d, generator=g)
        obs[t0:t1, :, p:] = torch.exp(torch.randn(t1 - t0, T, p, 32, dtype=dt, device=d, generator=g) * 0.8 - 1.5)
    return obs.reshape(-1)


def main():
    bits = int(os.environ.get("BITS", "64"))
    T = int(os.environ.get("FRAMES", "100000"))
    tiles_list = [int(x) for x in os.environ.get("TILES", "16,64,128,192,256,296").split(",")]
    modes = [m for m in os.environ.get("MODES", "v2,old12").split(",") if m]      # v2 = time-split pipelined scan; oldN = summary/fold/re-run, N warps per SM
    dt = torch.float64 if bits == 64 else torch.float32
    d = torch.device("cuda", 0)
    p, n = 4, 3
    mx = max(tiles_list)
    obs = make_obs(dt, mx, T, p, d)
    rng = np.random.default_rng(0)
    C = np.linalg.qr(rng.normal(size=(mx * 32, p, n)))[0]
    filt = dev.alloc_records(mx * 32, T, dev.sigma(n), dt, d)
    ym = dev.alloc_records(mx * 32, T, p, dt, d)
    es = obs.element_size()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    reps = int(os.environ.get("REPS", "3"))
    for tiles in tiles_list:
        B = tiles * 32
        prm = dev.TrackParams.from_numpy(np.zeros(n), np.eye(n), C[:B], np.eye(n), 1e-3 * np.eye(n), dtype=dt, device=d)
        tf = B * T
        f = b = 0.0
        os.environ["EKSB_ROUTE"] = "seq"
        for r in range(reps + 1):
            ev[0].record()
            dev.kf_forward(obs, prm, B, T, filt=filt)
            ev[1].record()
            dev.rts_backward(filt, prm, B, T, inplace=True, ym=ym)
            ev[2].record()
            torch.cuda.synchronize()
            if r:
                f += ev[0].elapsed_time(ev[1]) / reps; b += ev[1].elapsed_time(ev[2]) / reps
        os.environ.pop("EKSB_ROUTE", None)
        print(f"f{bits} tiles={tiles:4d} T={T} pipe   fwd {f:7.2f} bwd {b:7.2f} total {f+b:7.2f} ms -> {tf/(f+b)*1e-6:6.2f} G tf/s "
              f"({tf*(3*p+3*dev.sigma(n))*es/(f+b)*1e-6:6.0f} GB/s)", flush=True)
        for mode in modes:
            os.environ.pop("EKSB_SCAN_WARPS_PER_SM", None)
            if mode == "v2":
                os.environ["EKSB_SCAN_V2"] = "1"
            else:
                os.environ["EKSB_SCAN_V2"] = "0"
                os.environ["EKSB_SCAN_WARPS_PER_SM"] = mode[3:]
            ws_bytes = dev.lib().eksb_scan_workspace_bytes(B, T, n, 0, es)
            ws = torch.empty(ws_bytes // es + 1, dtype=dt, device=d)
            f = tot = 0.0
            for r in range(reps + 1):
                ev[0].record()
                dev.kf_rts_scan(obs, prm, B, T, smooth=False, filt=filt, workspace=ws)
                ev[1].record()
                out = dev.kf_rts_scan(obs, prm, B, T, smooth=True, inplace=True, filt=filt, ym=ym, workspace=ws)
                ev[2].record()
                torch.cuda.synchronize()
                if r:
                    f += ev[0].elapsed_time(ev[1]) / reps; tot += ev[1].elapsed_time(ev[2]) / reps
            assert int(out["status"].sum()) == 0
            print(f"f{bits} tiles={tiles:4d} T={T} {mode:6s} fwd {f:7.2f} bwd {tot-f:7.2f} total {tot:7.2f} ms -> {tf/tot*1e-6:6.2f} G tf/s "
                  f"({tf*(3*p+3*dev.sigma(n))*es/tot*1e-6:6.0f} GB/s) chunk={out['chunk']}", flush=True)
            del ws
        os.environ.pop("EKSB_SCAN_V2", None); os.environ.pop("EKSB_SCAN_WARPS_PER_SM", None)
    print("done")


if __name__ == "__main__":
    main()
