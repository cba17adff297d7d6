#!/usr/bin/env python
"""bench.py -- keypoint-frames/s smoothed (Kalman filter + RTS) on B200, against the HBM roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...   (N > 1)

One "step" = one pass of the hot path (eksb_kf_forward_ws + eksb_rts_backward_ws through the C ABI) over
BASELINE.json configs[3]: 16 384 synthetic tracks x 100 000 frames (5-model ensemble, 2 views -> p = 4
observations, n = 3 latents), float64 like the reference.  The 16 384 tracks are a FIXED problem sharded over
the N GPUs (contiguous blocks, selection.shard_bounds; strong scaling, no data-path collective): every rank
keeps the obs records of its whole shard resident in HBM and smooths them in groups of tracks whose output
buffers (filtered/smoothed records + C ms) fit beside them -- at N = 1 that is 105 GB of inputs plus the
buffers of one group.  The library routes a group with too few track tiles for the sequential kernels through
the parallel-in-time scan (the 2 048-track shard of N = 8).  Also in the line: parity checks of the bench batch
against the oracle, configs[4] (64 tracks x 10 M frames: scan filter+RTS and the block-tridiagonal solve),
the smoothing-parameter grid search with its one collective, the K1/K8 stages and the end-to-end host API.
Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "keypoint-frames/s smoothed (filter+RTS)"
UNIT = "keypoint-frames/s"
P_OBS, N_LAT, M_ENS = 4, 3, 5
SIGMA = N_LAT + N_LAT * (N_LAT + 1) // 2


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32"])
    ap.add_argument("--tracks", type=int, default=16384, help="TOTAL tracks of the job, sharded over the GPUs (configs[3]: 16384)")
    ap.add_argument("--group", type=int, default=0, help="tracks per group (default: the largest whose buffers fit)")
    ap.add_argument("--frames", type=int, default=100_000)
    ap.add_argument("--e2e-tracks", type=int, default=4096, help="TOTAL tracks of the end-to-end run, sharded over the GPUs")
    ap.add_argument("--e2e-block", type=int, default=0, help="tracks per streamed block of the e2e run (0 = the API's default sizing)")
    ap.add_argument("--c5-tracks", type=int, default=64)
    ap.add_argument("--c5-frames", type=int, default=10_000_000)
    ap.add_argument("--no-c5", action="store_true", help="skip configs[4] (64 x 10 M)")
    ap.add_argument("--no-grid", action="store_true", help="skip the smoothing-parameter grid search")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle checks of the bench batch")
    ap.add_argument("--cpu-seconds", type=float, default=20.0, help="target CPU-baseline sample time (per worker)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-setup", action="store_true", help="skip the K8 parameter-setup timing")
    ap.add_argument("--no-sustained", action="store_true", help="skip the sustained device-copy context measurement")
    return ap.parse_args()


# ----------------------------------------------------------------------------------------------
# synthetic workload: random-walk latents, per-track orthonormal C, log-normal
# per-frame noise scale with 1 % outlier frames, 5 ensemble members.
# ----------------------------------------------------------------------------------------------
def model_params(B, seed):
    rng = np.random.default_rng(seed)
    C = np.linalg.qr(rng.normal(size=(B, P_OBS, N_LAT)))[0]
    return dict(m0=np.zeros(N_LAT), S0=np.eye(N_LAT), C=C, A=np.eye(N_LAT), Q=1e-4 * np.eye(N_LAT))


def synth_preds_block(torch, C_blk, T, dtype, device, gen):
    """-> preds [M][T][nb*p] for one block of tracks (the raw 'models x frames x keypoints' tensor)."""
    nb = C_blk.shape[0]
    x = torch.randn(nb, T, N_LAT, dtype=dtype, device=device, generator=gen).mul_(0.1).cumsum_(1)
    clean = torch.einsum("bpn,btn->tbp", C_blk, x)                       # (T, nb, p)
    del x
    sd = torch.randn(T, nb, P_OBS, dtype=dtype, device=device, generator=gen).mul_(0.8).add_(float(np.log(0.5))).exp_()
    outl = torch.rand(T, nb, 1, dtype=dtype, device=device, generator=gen) < 0.01
    sd = torch.where(outl, sd * 30.0, sd)
    preds = torch.randn(M_ENS, T, nb, P_OBS, dtype=dtype, device=device, generator=gen)
    preds.mul_(sd.unsqueeze(0)).add_(clean.unsqueeze(0))
    return preds.reshape(M_ENS, T, nb * P_OBS)


class ClockSampler:
    """SM clock / power / throttle reasons sampled DURING the timed region (B200_PROFILING.md) through NVML --
    the source nvidia-smi reads -- from a thread of this process (no start-up delay, so short timed regions are
    still covered); falls back to an `nvidia-smi -lms` child process if pynvml is unavailable."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        self.rows, self.proc, self.index, self.stop_flag, self.thread, self.nv = [], None, index, False, None, None

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[self.index]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else self.index
            h = nv.nvmlDeviceGetHandleByIndex(phys)
            self.nv = nv
            masks = [getattr(nv, n, 0) for n in ("nvmlClocksEventReasonHwSlowdown", "nvmlClocksEventReasonHwThermalSlowdown",
                                                  "nvmlClocksEventReasonSwThermalSlowdown", "nvmlClocksEventReasonSwPowerCap")]
            fallb = [getattr(nv, n, 0) for n in ("nvmlClocksThrottleReasonHwSlowdown", "nvmlClocksThrottleReasonHwThermalSlowdown",
                                                  "nvmlClocksThrottleReasonSwThermalSlowdown", "nvmlClocksThrottleReasonSwPowerCap")]
            masks = [m or f for m, f in zip(masks, fallb)]
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)

            def loop():
                while not self.stop_flag:
                    try:
                        sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                        pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                        try:
                            r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                        except Exception:
                            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                        self.rows.append([str(sm), str(mx), f"{pw:.2f}"] + ["Active" if (r & m) else "Not Active" for m in masks])
                    except Exception:
                        pass
                    time.sleep(0.004)
            self.thread = threading.Thread(target=loop, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nv = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.thread is not None:
            self.stop_flag = True
            self.thread.join(timeout=1.0)
        elif self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
        else:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock sampling unavailable"]}
        rows = [r for r in self.rows if len(r) >= 7]
        sm = [float(r[0]) for r in rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in rows if r[1].replace(".", "").isdigit()]
        reasons = sorted({self.NAMES[i] for r in rows for i in range(4) if r[3 + i].lower().startswith("active")})
        pw = [float(r[2]) for r in rows if r[2].replace(".", "").isdigit()]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": reasons,
                "source": "nvml" if self.thread is not None else "nvidia-smi"}


def bind_to_gpu_numa_node(local_index):
    """Pin this process to the CPUs of the NUMA node its GPU hangs off (first-touch then places the pinned
    staging buffers of the e2e path on that node; with 8 ranks on one box this is what keeps eight PCIe
    streams from sharing one socket's memory controller).  Best effort; returns the node or None."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = int(vis.split(",")[local_index]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else local_index
        bus = nv.nvmlDeviceGetPciInfo(nv.nvmlDeviceGetHandleByIndex(phys)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        bus = bus.lower()
        if len(bus.split(":")[0]) == 8:
            bus = bus[4:]
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return node
    except Exception:
        return None
    return None


# ----------------------------------------------------------------------------------------------
# CPU arm: the reference's algorithm (numpy restatement in oracle/, pinned against the reference's
# own goldens) on the box's host cores.  The reference is pure Python and cannot travel to the GPU
# box, so kind = "port".  Bounded sample of the same workload: `tracks` tracks x `frames` frames
# per worker process, one worker per host core.
# ----------------------------------------------------------------------------------------------
def _cpu_worker(args):
    seed, tracks, frames = args
    try:
        from threadpoolctl import threadpool_limits
        limiter = threadpool_limits(limits=1)
    except Exception:
        limiter = None
    from oracle import restated as o
    rng = np.random.default_rng(seed)
    prm = model_params(tracks, seed)
    done = 0
    tc = 0.0
    for b in range(tracks):
        x = np.cumsum(rng.normal(size=(frames, N_LAT)) * 0.1, axis=0)
        sd = np.exp(rng.normal(np.log(0.5), 0.8, size=(frames, P_OBS)))
        preds = (x @ prm["C"][b].T)[None] + sd[None] * rng.normal(size=(M_ENS, frames, P_OBS))
        y, ev = o.ensemble_arrays(preds, "median")
        t1 = time.perf_counter()
        R = np.eye(P_OBS)
        mf, Vf, S = o.filtering_pass(y, prm["m0"], prm["S0"], prm["C"][b], R, prm["A"], prm["Q"], ev)
        ms, Vs, _ = o.smooth_backward(y, mf, Vf, S, prm["A"], prm["Q"], prm["C"][b])
        ym = ms @ prm["C"][b].T
        tc += time.perf_counter() - t1          # filter + RTS + back-projection only (the metric's scope)
        done += frames
    return done, tc, float(ym[-1, 0])


def cpu_arm(seconds, frames=20_000):
    """-> (track-frames/s aggregate, cores, sample description, wall seconds)."""
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    tracks = max(1, int(round(seconds * 11_000 / frames)))
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(cores) as pool:
        res = pool.map(_cpu_worker, [(1000 + i, tracks, frames) for i in range(cores)])
    wall = time.perf_counter() - t0
    total = sum(r[0] for r in res)
    busy = max(r[1] for r in res)               # slowest worker's filter+RTS time (input synthesis excluded)
    sample = (f"{cores} worker processes x {tracks} tracks x {frames} frames (p={P_OBS}, n={N_LAT}, M={M_ENS}), "
              f"numpy float64 filtering_pass+smooth_backward (oracle/restated.py), OpenBLAS 1 thread/worker")
    return total / busy, cores, sample, wall


def reference_main(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals = []
    for _ in range(a.warmup):
        cpu_arm(min(a.cpu_seconds, 2.0))
    t_tot = 0.0
    for _ in range(a.steps):
        v, cores, sample, wall = cpu_arm(min(a.cpu_seconds, 12.0))      # K bounded samples: the whole arm stays within minutes
        vals.append(v)
        t_tot += wall
    value = float(np.mean(vals))
    tracks = a.tracks
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * t_tot / max(1, a.steps), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"synthetic {tracks} tracks x {a.frames} frames, 5-model ensemble, 2 views (p=4, n=3) "
                               "[BASELINE configs[3]]; CPU arm timed on a bounded sample of it",
                   "host_cores": cores},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------
def measured_peak():
    try:
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(pk["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def committed_traffic(dtype, launch_tracks):
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture (profiles/traffic.json,
    tools/collect_profiles_r2.py), if that capture is of launches of the same size."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[dtype]
        return t["dram_bytes"] if t.get("launch_tracks") == launch_tracks else None
    except Exception:
        return None


def _es(dt_name):
    return 8 if dt_name == "f64" else 4


def rel_records(got, want, T, rows):
    """Largest error between two track-tile record buffers, every (track, component) scaled by its own largest
    magnitude over time (tests/conftest.py rel); evaluated in slices of frames to bound temporaries."""
    import torch
    g = got.view(-1, T, rows, 32)
    w = want.view(-1, T, rows, 32)
    step = max(1, (1 << 25) // (rows * 32))
    worst = 0.0
    for i in range(g.shape[0]):
        scale = torch.zeros(1, rows, 32, dtype=w.dtype, device=w.device)
        for t0 in range(0, T, step):
            scale = torch.maximum(scale, w[i, t0:t0 + step].abs().amax(dim=0, keepdim=True))
        scale = scale.clamp_min(max(float(scale.max()) * 1e-6, 1e-300))
        for t0 in range(0, T, step):
            worst = max(worst, float(((g[i, t0:t0 + step] - w[i, t0:t0 + step]).abs() / scale).max()))
    return worst


def track_rows(buf, T, rows, b, t0, t1, r0, r1):
    """rows [r0,r1) of frames [t0,t1) of track b out of a track-tile record buffer -> host float64 (t1-t0, r1-r0)."""
    return buf.view(-1, T, rows, 32)[b // 32, t0:t1, r0:r1, b % 32].double().cpu().numpy()


def oracle_parity(obs, par, T, dt, tracks, filt, ym, fwd, bwd, n_check=2000):
    """Oracle check of the bench batch itself.  For each listed track:
      * the filtered prefix mf, Vf[:n_check] depends on y, ev[:n_check] only -> compared with oracle.filtering_pass;
      * the smoothed ms, Vs and C ms of the LAST n_check frames -> compared with the oracle run on the last 2*n_check
        frames from a vague prior (filter and smoother forget their start: after n_check frames nothing of it is left).
    obs / filt / ym: the record buffers of ONE group of tracks (the bench's own), par["C"] that group's matrices;
    fwd() / bwd() run the bench's own forward / backward call on them.  Returns the largest per-component relative error."""
    import torch
    from oracle import restated as o
    sg = SIGMA
    fwd()
    torch.cuda.synchronize()
    iu = np.triu_indices(N_LAT)
    cast = (lambda a: a.astype(np.float32).astype(np.float64)) if dt == torch.float32 else (lambda a: a)
    worst = 0.0

    def upd(a, b):
        nonlocal worst
        scale = np.maximum(np.abs(b).max(axis=0, keepdims=True), max(np.abs(b).max() * 1e-6, 1e-300))
        worst = max(worst, float((np.abs(a - b) / scale).max()))

    n1 = min(n_check, T)
    pre = {}
    for b in tracks:
        y = track_rows(obs, T, 2 * P_OBS, b, 0, n1, 0, P_OBS)
        ev = track_rows(obs, T, 2 * P_OBS, b, 0, n1, P_OBS, 2 * P_OBS)
        C = cast(par["C"][b])
        mf, Vf, _ = o.filtering_pass(y, par["m0"], par["S0"], C, np.eye(P_OBS), par["A"], cast(par["Q"]), ev)
        got = track_rows(filt, T, sg, b, 0, n1, 0, sg)
        upd(got[:, :N_LAT], mf)
        upd(got[:, N_LAT:], Vf[:, iu[0], iu[1]])
    bwd()
    torch.cuda.synchronize()
    n2 = min(2 * n_check, T)
    for b in tracks:
        y = track_rows(obs, T, 2 * P_OBS, b, T - n2, T, 0, P_OBS)
        ev = track_rows(obs, T, 2 * P_OBS, b, T - n2, T, P_OBS, 2 * P_OBS)
        C = cast(par["C"][b])
        # vague prior centred on the least-squares latent of the first frame of the suffix
        m0 = np.linalg.lstsq(C, y[0], rcond=None)[0]
        mf, Vf, S = o.filtering_pass(y, m0, 1e3 * np.eye(N_LAT), C, np.eye(P_OBS), par["A"], cast(par["Q"]), ev)
        ms, Vs, _ = o.smooth_backward(y, mf, Vf, S, par["A"], cast(par["Q"]), C)
        k = n2 - min(n_check, n2 // 2)
        got = track_rows(filt, T, sg, b, T - n2 + k, T, 0, sg)
        upd(got[:, :N_LAT], ms[k:])
        upd(got[:, N_LAT:], Vs[k:][:, iu[0], iu[1]])
        upd(track_rows(ym, T, P_OBS, b, T - n2 + k, T, 0, P_OBS), ms[k:] @ C.T)
    return worst


def pcie_bandwidth(torch, d, nbytes=1 << 30, reps=4):
    """Pinned host <-> device copy bandwidth on this box, measured in-run (GB/s each way, one direction at a time)."""
    h = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    g = torch.empty(nbytes, dtype=torch.uint8, device=d)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    out = {}
    for name, (dst, src) in (("h2d", (g, h)), ("d2h", (h, g))):
        dst.copy_(src, non_blocking=True); torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            dst.copy_(src, non_blocking=True)
        e1.record(); torch.cuda.synchronize()
        out[name] = nbytes * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9
    del h, g
    return out


def gpu_main(a):
    import torch
    import torch.distributed as dist
    from eks_b200 import device as dev, batched, selection as sel, _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    d = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=d)
    numa_node = bind_to_gpu_numa_node(local) if world > 1 else None

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu:
        v, cores, sample, _ = cpu_arm(a.cpu_seconds)          # before any big pinned allocation / CUDA work
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}

    dt = torch.float64 if a.dtype == "f64" else torch.float32
    es = _es(a.dtype)
    T = a.frames
    lo, hi = sel.shard_bounds(a.tracks, rank, world)           # this rank's contiguous block of the fixed problem
    B = hi - lo
    par_all = model_params(a.tracks, 1234)                     # one parameter set for the whole job, sliced per shard
    par = dict(par_all, C=par_all["C"][lo:hi])
    launches = 0

    # ---- synthesis + K1 (ensemble, fused restack): preds [M][T][blk*p] -> obs records of the whole shard, block by block
    gen = torch.Generator(device=d); gen.manual_seed(1234 + rank)
    obs = dev.alloc_records(B, T, 2 * P_OBS, dt, d)
    C_dev = torch.as_tensor(par["C"], dtype=dt).to(d)
    Bs_setup = min(B, 8192)                                    # K8 is measured on the first <= 8192 tracks of the shard
    ws_setup = None if a.no_setup else dev.setup_workspace(Bs_setup, T, P_OBS, dt, d)
    mx_all = None if a.no_setup else dev.setup_max_var_view(ws_setup, Bs_setup, T, P_OBS, dt)
    blk = 128
    tile_elems = T * 2 * P_OBS * 32
    k1_ms, k1_tf = 0.0, 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for b0 in range(0, B, blk):
        nb = min(blk, B - b0)
        preds = synth_preds_block(torch, C_dev[b0:b0 + nb], T, dt, d, gen)
        out = obs[(b0 // 32) * tile_elems:((b0 + nb + 31) // 32) * tile_elems]
        mxb = None
        if mx_all is not None and b0 + nb <= Bs_setup:
            mxb = mx_all[(b0 // 32) * T * 32:((b0 + nb + 31) // 32) * T * 32]
        e0.record()
        dev.ensemble_tiled(preds, nb, P_OBS, "median", out=out, max_var=mxb)
        e1.record()
        torch.cuda.synchronize()
        if b0 > 0 and mxb is not None:                         # skip the first (cold) launch; same variant throughout
            k1_ms += e0.elapsed_time(e1); k1_tf += nb * T
        launches += 1
        del preds
    del out, mxb                                               # views of obs: they would keep it alive past `del obs`
    torch.cuda.empty_cache()
    peak, peak_src = measured_peak()

    # ---- K8 (parameter setup on the device) on the first Bs_setup tracks, before the output buffers exist (its
    # workspace is freed again: at N = 1 the 105 GB of inputs leave room for little else).  The centring of
    # multiview_pca_smoother.py:702-705 is fused into the forward filter (per-track observation offset), so the setup
    # leaves the records alone; a forward pass WITH the offsets is timed after the main region.  Not part of `value`.
    setup, means_s = None, None
    if not a.no_setup:
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dev.pca_setup(obs, Bs_setup, T, P_OBS, 25.0, 0.01, center=False, workspace=ws_setup, mx_ready=True)     # warm-up
        torch.cuda.synchronize()
        s0.record()
        prm_s, means_s, st_setup = dev.pca_setup(obs, Bs_setup, T, P_OBS, 25.0, 0.01, center=False, workspace=ws_setup, mx_ready=True)
        s1.record(); torch.cuda.synchronize()
        sms = s0.elapsed_time(s1)
        sb = (P_OBS + 1) * es                                      # y and max_vars (written by the ensemble kernel), read once
        tfs = Bs_setup * T
        setup = {"kernel": "pca_setup (percentile select + masked moments + PCA; max_vars from ensemble_tiled, centring fused into kf_forward)",
                 "tracks": Bs_setup, "ms": sms, "keypoint_frames_per_s": tfs / (sms * 1e-3),
                 "achieved_gbs": tfs * sb / (sms * 1e-3) / 1e9, "frac": tfs * sb / (sms * 1e-3) / 1e9 / peak,
                 "failed_tracks": int(st_setup.sum())}
        try:                                                       # committed ncu capture of one call of this shape (tools/gpu_ncu_r2.sh)
            tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("pca_setup_" + a.dtype)
            if tr and tr["algorithmic"] == tfs * sb:
                setup["dram_bytes_ncu"] = tr["core_dram_bytes"]
                setup["dram_bytes_over_algorithmic"] = tr["core_dram_bytes"] / (tfs * sb)
        except Exception:
            pass
        launches += 48
        del prm_s, ws_setup, mx_all
        torch.cuda.empty_cache()

    # ---- groups: the largest number of tracks whose output buffers fit beside the resident inputs
    per_track = T * (SIGMA + P_OBS) * es
    free_b, total_b = torch.cuda.mem_get_info(d)
    margin = 256 << 20                                         # headroom for the small tensors allocated after this
    filt = ym = None
    cands = [min(B, a.group)] if a.group else [c for c in (B, 8192, 4096, 2048, 1024, 512) if c <= B]
    for G in cands:                                            # try the allocation itself: at N = 1 the 8192-track
        if G * per_track + margin > free_b and G != cands[-1]:  # group is within a GiB of the board's capacity
            continue
        try:
            filt = dev.alloc_records(G, T, SIGMA, dt, d)
            ym = dev.alloc_records(G, T, P_OBS, dt, d)
            break
        except torch.OutOfMemoryError:
            filt = ym = None
            torch.cuda.empty_cache()
    assert filt is not None, "no group size fits beside the resident inputs"
    print(f"[bench] rank {rank}: shard {B} tracks, free after inputs {free_b / 2**30:.2f} GiB of {total_b / 2**30:.2f}, "
          f"group {G} tracks ({G * per_track / 2**30:.2f} GiB of buffers)", file=sys.stderr, flush=True)
    n_groups = (B + G - 1) // G
    prm_g, obs_g, ws_g = [], [], []
    for gi in range(n_groups):
        g0, g1 = gi * G, min(B, (gi + 1) * G)
        prm_g.append(dev.TrackParams.from_numpy(par["m0"], par["S0"], par["C"][g0:g1], par["A"], par["Q"], dtype=dt, device=d))
        obs_g.append(obs[(g0 // 32) * tile_elems:((g1 + 31) // 32) * tile_elems])
    Gmax = min(G, B)
    status = torch.zeros(B, dtype=torch.int32, device=d)
    ws = dev.route_workspace(Gmax, T, N_LAT, P_OBS, dt, d)     # None: this group size always runs the sequential kernels
    last = B - (n_groups - 1) * G
    if last != Gmax:
        ws_last = dev.route_workspace(last, T, N_LAT, P_OBS, dt, d)
        if ws_last is not None and (ws is None or ws_last.numel() > ws.numel()):
            ws = ws_last
    routed = [dev.routes_to_scan(min(B, (gi + 1) * G) - gi * G, T, N_LAT, P_OBS, dt) and ws is not None for gi in range(n_groups)]

    def fwd(gi):
        nb = min(B, (gi + 1) * G) - gi * G
        dev.kf_forward(obs_g[gi], prm_g[gi], nb, T, filt=filt, status=status[gi * G:gi * G + nb], workspace=ws)

    def bwd(gi):
        nb = min(B, (gi + 1) * G) - gi * G
        dev.rts_backward(filt, prm_g[gi], nb, T, inplace=True, ym=ym, status=status[gi * G:gi * G + nb], workspace=ws)

    for _ in range(a.warmup):
        for gi in range(n_groups):
            fwd(gi); bwd(gi)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = ClockSampler(local); clocks.start()
    ev = [[[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(n_groups)] for _ in range(a.steps)]
    for k in range(a.steps):
        for gi in range(n_groups):
            ev[k][gi][0].record()
            fwd(gi)
            ev[k][gi][1].record()
            bwd(gi)
            ev[k][gi][2].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clk = clocks.stop()
    total_ms = ev[0][0][0].elapsed_time(ev[-1][-1][2])
    fwd_ms = float(np.mean([sum(g[0].elapsed_time(g[1]) for g in st) for st in ev]))      # per step, all groups
    bwd_ms = float(np.mean([sum(g[1].elapsed_time(g[2]) for g in st) for st in ev]))
    assert int(status.sum()) == 0, "non-positive pivot / variance in the synthetic workload"
    assert bool(torch.isfinite(ym[:: 4099]).all())
    if world > 1:
        tmax = torch.tensor([total_ms], dtype=torch.float64, device=d)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        total_ms = float(tmax.item())
    value = a.tracks * T * a.steps / (total_ms * 1e-3)
    launches += (a.steps + a.warmup) * sum(10 if r else 2 for r in routed)

    # ---- roofline of the dominant kernel (algorithmic bytes per track-frame: DESIGN.md section 4).  A "launch" here is
    # one call of the entry point on one group; when the library routes the group through the parallel-in-time scan the
    # call is a warm-up kernel (K frames per chunk) + the pipelined kernel over (tile, chunk) virtual tiles (> 90 % of
    # it, see profiles/) + three fallback kernels that exit at once.
    tf_group = Gmax * T
    n_calls = n_groups
    kern = {"kf_forward": {"ms_per_step": fwd_ms, "ms_per_launch": fwd_ms / n_calls, "bytes_per_tf": (2 * P_OBS + SIGMA) * es},
            "rts_backward": {"ms_per_step": bwd_ms, "ms_per_launch": bwd_ms / n_calls, "bytes_per_tf": (2 * SIGMA + P_OBS) * es}}
    for k in kern.values():
        k["achieved_gbs"] = B * T * k["bytes_per_tf"] / (k["ms_per_step"] * 1e-3) / 1e9
        k["frac"] = k["achieved_gbs"] / peak
    dom = max(kern, key=lambda k: kern[k]["ms_per_step"])
    any_routed, all_routed = any(routed), all(routed)
    route = ("time-split pipelined scan (warm-up + pipelined kernels over (tile, chunk) virtual tiles)" if all_routed
             else ("mixed" if any_routed else "sequential warp-specialised kernels"))
    kname = {"kf_forward": ("eksb_kf_forward_ws -> " + ("scan_fwd_warm_kernel + kf_forward_pipe_kernel over virtual tiles" if all_routed else "kf_forward_pipe_kernel")),
             "rts_backward": ("eksb_rts_backward_ws -> " + ("scan_bwd_warm_kernel + rts_backward_pipe_kernel over virtual tiles" if all_routed else "rts_backward_pipe_kernel"))}
    traffic = committed_traffic(a.dtype, Gmax) if (dom == "rts_backward" and T == 100_000) else None
    step_b = B * T * (3 * P_OBS + 3 * SIGMA) * es
    roof = {"bound": "hbm", "kernel": kname[dom], "achieved": kern[dom]["achieved_gbs"], "peak": peak, "unit": "GB/s",
            "frac": kern[dom]["frac"], "traffic": traffic, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": tf_group * kern[dom]["bytes_per_tf"], "launch_tracks": Gmax,
            "ms_per_launch": kern[dom]["ms_per_launch"],
            "step_achieved": step_b / ((fwd_ms + bwd_ms) * 1e-3) / 1e9,
            "step_frac": step_b / ((fwd_ms + bwd_ms) * 1e-3) / 1e9 / peak,
            "kernels": kern, "route": route}

    # ---- parity of the bench batch itself: oracle on a few tracks' prefixes / suffixes, and (when routed) the
    # parallel-in-time result against the sequential kernels on the same group
    parity = None
    if not a.no_parity:
        tol = 1e-9 if a.dtype == "f64" else 1e-3
        worst, checked = 0.0, []
        for gi in sorted({0, n_groups - 1}):
            nb = min(B, (gi + 1) * G) - gi * G
            tr = sorted({0, nb // 3, nb - 1})
            parg = dict(par, C=par["C"][gi * G:gi * G + nb])
            worst = max(worst, oracle_parity(obs_g[gi], parg, T, dt, tr, filt, ym, lambda gi=gi: fwd(gi), lambda gi=gi: bwd(gi)))
            checked += [lo + gi * G + b for b in tr]
            launches += 10 if routed[gi] else 2
        parity = {"max_rel": worst, "tol": tol, "tracks": checked, "frames_each": "first 2000 filtered + last 2000 smoothed",
                  "oracle": "oracle/restated.py filtering_pass + smooth_backward", "metric": "per-component max over time"}
        if any_routed and free_b > 2 * Gmax * per_track + margin:
            gi = routed.index(True)
            nb = min(B, (gi + 1) * G) - gi * G
            fwd(gi); bwd(gi)                                        # routed result in filt / ym
            f2 = dev.alloc_records(nb, T, SIGMA, dt, d); y2 = dev.alloc_records(nb, T, P_OBS, dt, d)
            st2 = torch.zeros(nb, dtype=torch.int32, device=d)
            dev.kf_forward(obs_g[gi], prm_g[gi], nb, T, filt=f2, status=st2)          # no workspace: sequential kernels
            dev.rts_backward(f2, prm_g[gi], nb, T, inplace=True, ym=y2, status=st2)
            torch.cuda.synchronize()
            r1 = rel_records(filt[:f2.numel()], f2, T, SIGMA)
            r2 = rel_records(ym[:y2.numel()], y2, T, P_OBS)
            parity["scan_vs_sequential_max_rel"] = max(r1, r2)
            launches += 8
            del f2, y2
            torch.cuda.empty_cache()
        ok = worst <= tol and parity.get("scan_vs_sequential_max_rel", 0.0) <= tol
        parity["ok"] = bool(ok)
        assert ok, f"parity check of the bench batch failed: {parity}"

    # ---- forward pass with the camera means as observation offset (the session pipeline's variant), group 0
    fwd_off_ms = None
    if means_s is not None:
        nb0 = min(G, B, Bs_setup)
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for rep in range(2):
            s0.record()
            dev.kf_forward(obs_g[0], prm_g[0], nb0, T, filt=filt, status=status[:nb0], y_offset=means_s[:nb0].contiguous(), workspace=ws)
            s1.record(); torch.cuda.synchronize()
        fwd_off_ms = s0.elapsed_time(s1) * (B / nb0)             # scaled to the whole shard, like fwd_ms
        launches += 2 * (5 if routed[0] else 1)
    k1 = None
    if k1_ms > 0:
        k1b = (M_ENS * P_OBS + 2 * P_OBS + 1) * es                 # + the max_vars row handed to the setup
        k1 = {"kernel": "ensemble_tiled", "keypoint_frames_per_s": k1_tf / (k1_ms * 1e-3),
              "achieved_gbs": k1_tf * k1b / (k1_ms * 1e-3) / 1e9, "frac": k1_tf * k1b / (k1_ms * 1e-3) / 1e9 / peak}
    session = None
    if k1 and setup and fwd_off_ms:
        per_tf = 1.0 / k1["keypoint_frames_per_s"] + 1.0 / setup["keypoint_frames_per_s"] + (fwd_off_ms + bwd_ms) * 1e-3 / (B * T)
        session = {"keypoint_frames_per_s": world * 1.0 / per_tf, "stages": "ensemble_tiled + pca_setup + kf_forward (y offset = camera means) + rts_backward",
                   "kf_forward_with_offset_ms": fwd_off_ms,
                   "note": "raw [models x frames x keypoints] predictions in HBM -> smoothed records, sum of the measured stage times (all GPUs)"}

    # ---- smoothing-parameter grid search on this rank's shard (first group) and the path's ONE collective: the
    # all-reduce of the (n_grid,) NLL vector (replaces the CLI flag of scripts/multicam_example.py:38-43)
    grid = None
    if not a.no_grid:
        s_grid = np.logspace(-2, 1.75, 16)
        nbg = min(G, B)
        g0e, g1e, g2e = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        sel.nll_grid(obs_g[0], prm_g[0], min(nbg, 256), T, s_grid[:2])                  # warm-up (module load)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        g0e.record()
        nll, _ = sel.nll_grid(obs_g[0], prm_g[0], nbg, T, s_grid)
        local_sums = sel.nll_sums(nll)
        g1e.record()
        tot = local_sums.clone()
        sel.allreduce_sum_(tot)
        g2e.record()
        torch.cuda.synchronize()
        best = int(torch.argmin(tot).item())
        grid = {"n_grid": len(s_grid), "tracks_per_gpu": nbg, "ms": g0e.elapsed_time(g1e), "allreduce_us": 1e3 * g1e.elapsed_time(g2e),
                "best_s": float(s_grid[best]), "collective": "nccl all_reduce(SUM) of 16 float64" if world > 1 else "none (1 GPU)",
                "nll_evaluations_per_s": world * nbg * T * len(s_grid) / (g0e.elapsed_time(g1e) * 1e-3)}
        launches += 4
        if world > 1:
            gathered = [torch.empty_like(local_sums) for _ in range(world)]
            dist.all_gather(gathered, local_sums)
            want = torch.stack(gathered).sum(0)
            bests = [torch.empty(1, dtype=torch.int64, device=d) for _ in range(world)]
            dist.all_gather(bests, torch.tensor([best], dtype=torch.int64, device=d))
            agree = all(int(b.item()) == best for b in bests)
            close = bool(torch.allclose(tot, want, rtol=1e-12, atol=0.0))
            grid["all_ranks_pick_same_s"] = agree
            grid["allreduce_equals_sum_of_gathered"] = close
            assert agree and close, f"grid search collective check failed: {grid}"

    # ---- free the main buffers; context for the roofline: what a plain device copy SUSTAINS on this board right now
    del filt, ym, obs_g, prm_g, ws
    torch.cuda.empty_cache()
    sustained = None
    if rank == 0 and not a.no_sustained:
        nbytes = 4 << 30
        src = torch.empty(nbytes, dtype=torch.uint8, device=d); dst = torch.empty_like(src)
        src.zero_(); dst.zero_()
        sc = ClockSampler(local); sc.start()
        for _ in range(20):
            dst.copy_(src)
        torch.cuda.synchronize()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 400
        c0.record()
        for _ in range(reps):
            dst.copy_(src)
        c1.record(); torch.cuda.synchronize()
        sclk = sc.stop()
        sustained = {"gbs": 2 * nbytes * reps / (c0.elapsed_time(c1) * 1e-3) / 1e9, "seconds": c0.elapsed_time(c1) * 1e-3,
                     "how": "dst.copy_(src), 4 GiB buffers, 400 back-to-back copies after 20 warm-up copies",
                     "sm_mhz": sclk["sm_mhz"], "reasons": sclk["reasons"]}
        del src, dst
    roof["sustained_copy"] = sustained
    if sustained:
        roof["step_frac_of_sustained_copy"] = roof["step_achieved"] / sustained["gbs"]
    del obs
    torch.cuda.empty_cache()

    # ---- configs[4]: few very long tracks (64 x 10 M frames), rank 0 of a 1-GPU run only
    c5 = None
    if rank == 0 and world == 1 and not a.no_c5:
        c5 = config5(a, torch, dev, d, dt, es, peak)
        launches += c5.pop("_launches")

    # ---- e2e: public host API, H2D + D2H inside the timed region; a FIXED batch sharded over the ranks
    e2e = None
    if not a.no_e2e:
        e2e_total = a.e2e_tracks
        try:                                                    # three pinned (B,T,p) arrays per rank must fit the host
            import psutil
            avail = psutil.virtual_memory().available
            while e2e_total > 256 * world and 3 * e2e_total * T * P_OBS * es > 0.6 * avail:
                e2e_total //= 2
        except Exception:
            pass
        a.e2e_tracks = e2e_total
        elo, ehi = sel.shard_bounds(a.e2e_tracks, rank, world)
        Be = ehi - elo
        pe = model_params(a.e2e_tracks, 99)
        pe["C"] = pe["C"][elo:ehi]
        y_h = torch.empty((Be, T, P_OBS), dtype=dt, pin_memory=True)
        ev_h = torch.empty((Be, T, P_OBS), dtype=dt, pin_memory=True)
        for b0 in range(0, Be, 128):                            # synthesise on the device, park in pinned host memory
            nb = min(128, Be - b0)
            Cb = torch.as_tensor(pe["C"][b0:b0 + nb], dtype=dt, device=d)
            pr = synth_preds_block(torch, Cb, T, dt, d, gen)
            avg, var = dev.ensemble(pr, "median")
            y_h[b0:b0 + nb].copy_(avg.reshape(T, nb, P_OBS).permute(1, 0, 2))
            ev_h[b0:b0 + nb].copy_(var.reshape(T, nb, P_OBS).permute(1, 0, 2))
            del pr, avg, var
        torch.cuda.empty_cache()
        pcie = pcie_bandwidth(torch, d)
        outs = batched.alloc_outputs(Be, T, N_LAT, P_OBS, dt, pin=True, outputs=("ym",))
        cnt = {}
        call = lambda c=None: batched.smooth_tracks(y_h, ev_h, pe["m0"], pe["S0"], pe["C"], pe["A"], pe["Q"], dtype=dt,
                                                    device=d, block_tracks=a.e2e_block or None, out=outs, outputs=("ym",), counters=c)
        call()                                                  # warm-up (allocator, first-touch)
        e2e_steps = max(1, min(a.steps, 5))
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            call(cnt)
        torch.cuda.synchronize()
        dt_s = time.perf_counter() - t0
        if world > 1:
            tm = torch.tensor([dt_s], dtype=torch.float64, device=d)
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            dt_s = float(tm.item())
        assert int(outs["status"].sum()) == 0 and bool(torch.isfinite(outs["ym"][::97]).all())
        h2d, d2h = cnt["h2d_bytes"] // e2e_steps, cnt["d2h_bytes"] // e2e_steps
        step_s = dt_s / e2e_steps
        e2e = {"value": a.e2e_tracks * T * e2e_steps / dt_s, "unit": UNIT,
               "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": d2h * world,
               "tracks": a.e2e_tracks, "tracks_per_gpu": Be, "frames": T, "steps": e2e_steps, "numa_node": numa_node,
               "outputs": "ym = C ms (what every reference smoother consumes, multiview_pca_smoother.py:745,761-762)",
               "api": f"eks_b200.batched.smooth_tracks (pinned host in/out, {a.e2e_block or 'default-sized (~0.8 GB of input)'} track blocks, 3 streams per GPU)",
               "ms_per_step": 1e3 * step_s,
               "pcie": {"h2d_gbs_measured": pcie["h2d"], "d2h_gbs_measured": pcie["d2h"],
                        "h2d_gbs_in_run": h2d / step_s / 1e9, "d2h_gbs_in_run": d2h / step_s / 1e9,
                        "how": "1 GiB pinned <-> device copies, one direction at a time, this rank"},
               "host_dma_gbs_all_gpus": (h2d + d2h) * world / step_s / 1e9,     # what the host moved per second for all ranks
               "pcie_frac": max(h2d / step_s / 1e9 / pcie["h2d"], d2h / step_s / 1e9 / pcie["d2h"]),
               "binding_direction": "h2d" if h2d / pcie["h2d"] >= d2h / pcie["d2h"] else "d2h"}
        launches += cnt["launches"]

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": total_ms / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": a.dtype, "data": "synthetic",
            "config": {"workload": f"synthetic {a.tracks} tracks x {T} frames, 5-model ensemble, 2 views (p=4, n=3) "
                                   f"[BASELINE configs[3]], {a.dtype}, sharded over {world} GPU(s)",
                       "tracks_total": a.tracks, "tracks_per_gpu": B, "frames": T, "group_tracks": Gmax, "groups_per_gpu": n_groups,
                       "route": route,
                       "parallelism": f"fixed problem, contiguous track shards x{world}, no data-path collective",
                       "resident": f"obs records of the whole shard ({B * T * 2 * P_OBS * es / 1e9:.1f} GB) + output buffers of one group "
                                   f"({Gmax * per_track / 1e9:.1f} GB) of {total_b / 1e9:.1f} GB",
                       "l2": "inputs (obs records) exceed L2 by >100x; no flush needed",
                       "smoothing": "in place (smth overwrites filt), ym = C ms written"},
            "roofline": roof, "parity_check": parity, "config5": c5, "grid_search": grid,
            "ensemble_kernel": k1, "setup_kernel": setup, "session_pipeline": session, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": launches, "clocks": clk,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def config5(a, torch, dev, d, dt, es, peak):
    """BASELINE configs[4] / SURVEY C5: 64 tracks x 10 M frames.  Inputs y, ev directly, same generative
    model, synthesised in time slices.  Timed: parallel-in-time filter + RTS (the routed entry points) and the
    block-tridiagonal solve; gated: scan == sequential kernels on the whole batch, oracle on prefixes / suffixes."""
    B, T = a.c5_tracks, a.c5_frames
    gen = torch.Generator(device=d); gen.manual_seed(555)
    par = model_params(B, 4321)
    prm = dev.TrackParams.from_numpy(par["m0"], par["S0"], par["C"], par["A"], par["Q"], dtype=dt, device=d)
    tiles = dev.num_tiles(B)
    obs = dev.alloc_records(B, T, 2 * P_OBS, dt, d)
    ov = obs.view(tiles, T, 2 * P_OBS, 32)
    ov[:, :, :P_OBS] = 0; ov[:, :, P_OBS:] = 1                 # padding lanes: y = 0, ev = 1
    Cd = torch.as_tensor(par["C"], dtype=dt).to(d)
    x0 = torch.zeros(B, 1, N_LAT, dtype=dt, device=d)
    Bp = tiles * 32
    step_t = 500_000
    for t0 in range(0, T, step_t):
        t1 = min(T, t0 + step_t)
        x = torch.randn(B, t1 - t0, N_LAT, dtype=dt, device=d, generator=gen).mul_(0.1).cumsum_(1).add_(x0)
        x0 = x[:, -1:, :].clone()
        sd = torch.randn(B, t1 - t0, P_OBS, dtype=dt, device=d, generator=gen).mul_(0.8).add_(float(np.log(0.5))).exp_()
        outl = torch.rand(B, t1 - t0, 1, dtype=dt, device=d, generator=gen) < 0.01
        sd = torch.where(outl, sd * 30.0, sd)
        ev = sd * sd / M_ENS
        y = torch.einsum("bpn,btn->btp", Cd, x) + ev.sqrt() * torch.randn(B, t1 - t0, P_OBS, dtype=dt, device=d, generator=gen)
        pad = lambda z: torch.cat([z, z[:1].expand(Bp - B, -1, -1)], 0) if Bp > B else z
        ov[:, t0:t1, :P_OBS] = pad(y).view(tiles, 32, t1 - t0, P_OBS).permute(0, 2, 3, 1)
        ov[:, t0:t1, P_OBS:] = pad(ev).view(tiles, 32, t1 - t0, P_OBS).permute(0, 2, 3, 1)
        del x, sd, outl, ev, y
    filt = dev.alloc_records(B, T, SIGMA, dt, d)
    ym = dev.alloc_records(B, T, P_OBS, dt, d)
    status = torch.zeros(B, dtype=torch.int32, device=d)
    ws = dev.route_workspace(B, T, N_LAT, P_OBS, dt, d)
    assert ws is not None, "configs[4] must route to the parallel-in-time kernels"

    def fwd():
        dev.kf_forward(obs, prm, B, T, filt=filt, status=status, workspace=ws)

    def bwd():
        dev.rts_backward(filt, prm, B, T, inplace=True, ym=ym, status=status, workspace=ws)

    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    reps, f_ms, b_ms = 3, 0.0, 0.0
    for r in range(reps + 1):
        e[0].record(); fwd(); e[1].record(); bwd(); e[2].record()
        torch.cuda.synchronize()
        if r:
            f_ms += e[0].elapsed_time(e[1]) / reps; b_ms += e[1].elapsed_time(e[2]) / reps
    assert int(status.sum()) == 0
    tf = B * T
    step_b = tf * (3 * P_OBS + 3 * SIGMA) * es
    out = {"workload": f"{B} tracks x {T} frames (p=4, n=3), {a.dtype}, 1 GPU", "route": "time-split pipelined scan (eksb_kf_forward_ws + eksb_rts_backward_ws: warm-up + pipelined kernels over virtual tiles)",
           "chunk_frames": int(dev.lib().eksb_scan_chunk_len(B, T, N_LAT, 0, es)),
           "filter_rts": {"ms": f_ms + b_ms, "fwd_ms": f_ms, "bwd_ms": b_ms, "keypoint_frames_per_s": tf / ((f_ms + b_ms) * 1e-3),
                          "algorithmic_gbs": step_b / ((f_ms + b_ms) * 1e-3) / 1e9, "frac": step_b / ((f_ms + b_ms) * 1e-3) / 1e9 / peak,
                          "bytes_per_track_frame": (3 * P_OBS + 3 * SIGMA) * es}}
    launches = 10 * (reps + 1)
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["config5_" + a.dtype]
        out["filter_rts"]["dram_bytes_ncu"] = t["dram_bytes"]
        out["filter_rts"]["dram_bytes_over_algorithmic"] = t["dram_bytes"] / step_b
    except Exception:
        pass
    # gate 1: oracle on prefixes / suffixes of 4 tracks (fresh forward + backward on the bench's buffers)
    if not a.no_parity:
        tol = 1e-9 if a.dtype == "f64" else 1e-3
        tr = sorted({0, B // 3, (2 * B) // 3, B - 1})
        worst = oracle_parity(obs, par, T, dt, tr, filt, ym, fwd, bwd)
        launches += 10
        # gate 2: the whole batch against the sequential kernels (one warp per tile: ~1 s per pass)
        f2 = dev.alloc_records(B, T, SIGMA, dt, d)
        st2 = torch.zeros(B, dtype=torch.int32, device=d)
        dev.kf_forward(obs, prm, B, T, filt=f2, status=st2)
        dev.rts_backward(f2, prm, B, T, inplace=True, want_ym=False, status=st2)
        torch.cuda.synchronize()
        r_seq = rel_records(filt, f2, T, SIGMA)
        launches += 2
        del f2
        torch.cuda.empty_cache()
        out["parity_check"] = {"max_rel": worst, "scan_vs_sequential_max_rel": r_seq, "tol": tol, "tracks": tr,
                               "ok": bool(worst <= tol and r_seq <= tol)}
        assert out["parity_check"]["ok"], f"configs[4] parity failed: {out['parity_check']}"
    # block-tridiagonal Newton step (A7) on the same batch: x == RTS mean, diagonal blocks of H^-1 == Vs
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    bt_ms = 0.0
    for r in range(3):
        e0.record()
        sol = dev.btd_solve(obs, prm, B, T, want_ym=False, status=status)
        e1.record(); torch.cuda.synchronize()
        if r:
            bt_ms += e0.elapsed_time(e1) / 2
        if r < 2:
            del sol
    btd_b = tf * (2 * P_OBS + 3 * SIGMA) * es                      # filter + RTS without the ym write
    out["btd_solve"] = {"ms": bt_ms, "keypoint_frames_per_s": tf / (bt_ms * 1e-3), "algorithmic_gbs": btd_b / (bt_ms * 1e-3) / 1e9,
                        "frac": btd_b / (bt_ms * 1e-3) / 1e9 / peak,
                        "note": "routed to the parallel-in-time filter + smoother elements (x == RTS mean, H^-1 blocks == Vs: SURVEY fact 0.8); "
                                "the block-Thomas kernels serve batches with >= 1 tile per SM"}
    if not a.no_parity:
        fwd(); bwd(); torch.cuda.synchronize()
        out["btd_solve"]["vs_filter_rts_max_rel"] = rel_records(sol["sol"], filt, T, SIGMA)
        launches += 10
    launches += 30
    del sol, obs, filt, ym
    torch.cuda.empty_cache()
    out["_launches"] = launches
    return out


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        reference_main(args)
    else:
        gpu_main(args)
