"""Copy the evidence of the last tools/gpu_ncu_r2.sh run from gpurun_out/ (scratch) into profiles/ (tracked):
ncu launch list + summary with the kernels' shares of the timed step, full-size DRAM bytes -> traffic.json
(read by bench.py for roofline.traffic and config5.dram_bytes_ncu), ncu --set full raw page."""
import collections, csv, json, os, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
tag = sys.argv[1] if len(sys.argv) > 1 else "r2"


def table(path):
    rows = list(csv.reader(open(path)))
    i = [k for k, r in enumerate(rows) if r and r[0] == "ID"][0]
    return rows[i], rows[i + 1:]


def short(name):
    return name.split("(")[0].replace("void ", "").replace("eksb::", "")[:110]


for src, dst in (("launches.csv", f"{tag}_launches.csv"), ("dram_full.csv", f"{tag}_dram_bytes_full_size.csv"),
                 ("dram_c5.csv", f"{tag}_dram_bytes_config5.csv")):
    shutil.copy(os.path.join(G, src), os.path.join(P, dst))

bench = json.loads(open(os.path.join(P, f"{tag}_bench_f64.json")).read().strip().splitlines()[-1])
T = bench["config"]["frames"]
Gt = bench["config"]["group_tracks"]

# ---- (2) DRAM bytes of the dominant kernels, per launch
h, rows = table(os.path.join(G, "dram_full.csv"))
kn, mn, mv = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value")
acc = collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows:
    if len(r) > mv:
        acc["fwd" if "kf_forward" in r[kn] else "bwd"][r[mn]].append(float(r[mv].replace(",", "")))
mean = lambda k, m: sum(acc[k][m]) / len(acc[k][m])
tf = Gt * T
traffic = {"_source": f"profiles/{tag}_dram_bytes_full_size.csv: ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum "
                      f"--clock-control none -k regex:pipe_kernel on `python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-c5 ...` "
                      f"(f64, {Gt}-track launches x {T} frames), mean per launch",
           "f64": {"kernel": "rts_backward_pipe_kernel", "launch_tracks": Gt,
                   "dram_bytes": int(mean("bwd", "dram__bytes_read.sum") + mean("bwd", "dram__bytes_write.sum")),
                   "read": int(mean("bwd", "dram__bytes_read.sum")), "write": int(mean("bwd", "dram__bytes_write.sum")),
                   "algorithmic": tf * 176, "ncu_duration_ms": mean("bwd", "gpu__time_duration.sum") / 1e6,
                   "kf_forward_pipe_kernel": {"dram_bytes": int(mean("fwd", "dram__bytes_read.sum") + mean("fwd", "dram__bytes_write.sum")),
                                              "read": int(mean("fwd", "dram__bytes_read.sum")), "write": int(mean("fwd", "dram__bytes_write.sum")),
                                              "algorithmic": tf * 136, "ncu_duration_ms": mean("fwd", "gpu__time_duration.sum") / 1e6}}}

# ---- (2b) f32: one launch per pass covers the whole 16 384-track batch
f32_csv = os.path.join(G, "dram_full_f32.csv")
if os.path.exists(f32_csv):
    shutil.copy(f32_csv, os.path.join(P, f"{tag}_dram_bytes_full_size_f32.csv"))
    try:
        b32 = json.loads(open(os.path.join(P, f"{tag}_bench_f32.json")).read().strip().splitlines()[-1])
        G32 = b32["config"]["group_tracks"]
    except Exception:
        G32 = 16384
    h, rows = table(f32_csv)
    kn, mn, mv = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value")
    acc = collections.defaultdict(lambda: collections.defaultdict(list))
    for r in rows:
        if len(r) > mv:
            acc["fwd" if "kf_forward" in r[kn] else "bwd"][r[mn]].append(float(r[mv].replace(",", "")))
    tf32 = G32 * T
    traffic["f32"] = {"_source": f"profiles/{tag}_dram_bytes_full_size_f32.csv (same command with --dtype f32, {G32}-track launches)",
                      "kernel": "rts_backward_pipe_kernel", "launch_tracks": G32,
                      "dram_bytes": int(mean("bwd", "dram__bytes_read.sum") + mean("bwd", "dram__bytes_write.sum")),
                      "read": int(mean("bwd", "dram__bytes_read.sum")), "write": int(mean("bwd", "dram__bytes_write.sum")),
                      "algorithmic": tf32 * 88, "ncu_duration_ms": mean("bwd", "gpu__time_duration.sum") / 1e6,
                      "kf_forward_pipe_kernel": {"dram_bytes": int(mean("fwd", "dram__bytes_read.sum") + mean("fwd", "dram__bytes_write.sum")),
                                                 "algorithmic": tf32 * 68, "ncu_duration_ms": mean("fwd", "gpu__time_duration.sum") / 1e6}}

# ---- (2c) pca_setup: every kernel of one call (wide-digit select, the default), tools/setup_time.py bench
setup_csv = os.path.join(G, "dram_setup.csv")
if os.path.exists(setup_csv):
    shutil.copy(setup_csv, os.path.join(P, f"{tag}_dram_bytes_setup.csv"))
    h, rows = table(setup_csv)
    kn, mn, mv, idc = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
    per = collections.OrderedDict()
    for r in rows:
        if len(r) > mv:
            per.setdefault(r[idc], {"name": short(r[kn])})[r[mn]] = float(r[mv].replace(",", ""))
    ls = list(per.values())
    starts = [i for i, l in enumerate(ls) if l["name"].startswith("setup_init_state_kernel")]
    call = ls[starts[2]:starts[3]]                          # third call of the wide path (the first two warm up)
    core = [l for l in call if not l["name"].startswith(("setup_mx_kernel", "center_rows_kernel"))]
    byt = lambda l: l.get("dram__bytes_read.sum", 0) + l.get("dram__bytes_write.sum", 0)
    traffic["pca_setup_f64"] = {"_source": f"profiles/{tag}_dram_bytes_setup.csv: one eksb_pca_setup call, 8192 tracks x 100000 frames (tools/setup_time.py bench); "
                                           "'core' = without setup_mx_kernel / center_rows_kernel, which the bench's pipeline does not run (max_vars from the ensemble "
                                           "kernel, centring fused into the forward filter)",
                                "core_dram_bytes": int(sum(byt(l) for l in core)), "core_ms": sum(l.get("gpu__time_duration.sum", 0) for l in core) / 1e6,
                                "algorithmic": 8192 * 100000 * 40,
                                "kernels": [{"kernel": l["name"][:50], "ms": l.get("gpu__time_duration.sum", 0) / 1e6, "dram_bytes": int(byt(l))}
                                            for l in call if l.get("gpu__time_duration.sum", 0) > 50e3]}

# ---- (3) configs[4]: all kernels of one routed forward + backward call (the LAST complete call in the capture)
h, rows = table(os.path.join(G, "dram_c5.csv"))
kn, mn, mv, idc = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
per_id = collections.OrderedDict()
for r in rows:
    if len(r) > mv:
        per_id.setdefault(r[idc], {"name": short(r[kn])})[r[mn]] = float(r[mv].replace(",", ""))
launches = list(per_id.values())
# a call = warm-up, pipe, 3 gated fallback kernels, forward then backward: take the last 10 launches that form one
names = [l["name"] for l in launches]
last_b = max(i for i, n in enumerate(names) if n.startswith("rts_backward_pipe_kernel<double, 3, 4, 0, 0, 1"))   # with the ym write
first_f = max(i for i, n in enumerate(names[:last_b]) if n.startswith("scan_fwd_warm"))
call = launches[first_f:last_b + 4]
c5_bytes = sum(l.get("dram__bytes_read.sum", 0) + l.get("dram__bytes_write.sum", 0) for l in call)
c5_ms = sum(l.get("gpu__time_duration.sum", 0) for l in call) / 1e6
traffic["config5_f64"] = {"_source": f"profiles/{tag}_dram_bytes_config5.csv: one eksb_kf_forward_ws + eksb_rts_backward_ws call on 64 tracks x 10 M frames "
                                     "(SCAN_TIME_BIG_ONLY=1 SCAN_TIME_F64_ONLY=1 python tools/scan_time.py), all kernels of the call summed",
                          "dram_bytes": int(c5_bytes), "algorithmic": 64 * 10_000_000 * 312, "ncu_duration_ms": c5_ms,
                          "kernels": [{"kernel": l["name"][:60], "ms": l.get("gpu__time_duration.sum", 0) / 1e6,
                                       "dram_bytes": int(l.get("dram__bytes_read.sum", 0) + l.get("dram__bytes_write.sum", 0))} for l in call]}
json.dump(traffic, open(os.path.join(P, "traffic.json"), "w"), indent=1)

# ---- (1) launch list summary
h, rows = table(os.path.join(G, "launches.csv"))
kn, mv = h.index("Kernel Name"), h.index("Metric Value")
agg = collections.OrderedDict()
for r in rows:
    if len(r) > mv:
        agg.setdefault(short(r[kn]), []).append(float(r[mv].replace(",", "")))
out = ["# ncu --metrics gpu__time_duration.sum --clock-control none -k <this library's kernels> python bench.py --steps 2 --warmup 3 --no-cpu",
       f"# (B200, f64, 16384 tracks x {T} frames in launches of {Gt} tracks; serialised cold-cache times: compare SHARES)",
       "# launches  total_ms  ms/launch  kernel"]
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    out.append(f"{len(v):6d} {sum(v) / 1e6:10.2f} {sum(v) / len(v) / 1e6:9.3f}  {k}")
# the timed step: per group one forward call (warm-up + pipe) and one backward call; durations of the full-size launches
def mean_big(prefix, frac=0.5):
    v = [x for k, vals in agg.items() if k.startswith(prefix) for x in vals]
    big = [x for x in v if x >= frac * max(v)]
    return sum(big) / len(big) / 1e6
fm, bm = mean_big("kf_forward_pipe_kernel<double, 3, 4, 0, 0"), mean_big("rts_backward_pipe_kernel<double, 3, 4, 0, 0, 1")
# warm-up launches of the timed step = the ones right before a full-size pipelined launch (the grid search and configs[4]
# launch the same kernels with other shapes)
def warm_before(pipe_prefix, warm_prefix):
    big = max(x for k, vals in agg.items() if k.startswith(pipe_prefix) for x in vals)
    last, got = None, []
    for r in rows:
        if len(r) <= mv:
            continue
        n = short(r[kn])
        if n.startswith(warm_prefix):
            last = float(r[mv].replace(",", ""))
        elif n.startswith(pipe_prefix):
            if last is not None and float(r[mv].replace(",", "")) >= 0.5 * big:
                got.append(last)
            last = None
    return sum(got) / len(got) / 1e6 if got else 0.0
wf = warm_before("kf_forward_pipe_kernel<double, 3, 4, 0, 0", "scan_fwd_warm")
wb = warm_before("rts_backward_pipe_kernel<double, 3, 4, 0, 0, 1", "scan_bwd_warm")
kf, kb = bench["roofline"]["kernels"]["kf_forward"]["ms_per_launch"], bench["roofline"]["kernels"]["rts_backward"]["ms_per_launch"]
tot = fm + bm + wf + wb
out += [f"# timed step, per {Gt}-track group: scan_fwd_warm {wf:.3f} + kf_forward_pipe {fm:.3f} + scan_bwd_warm {wb:.3f} + rts_backward_pipe {bm:.3f} ms "
        f"= shares {100 * wf / tot:.1f} / {100 * fm / tot:.1f} / {100 * wb / tot:.1f} / {100 * bm / tot:.1f} %  (forward call {100 * (wf + fm) / tot:.1f} %, backward call {100 * (wb + bm) / tot:.1f} %;",
        f"#   bench CUDA events, same code: forward call {kf:.2f} ms, backward call {kb:.2f} ms = {100 * kf / (kf + kb):.1f} / {100 * kb / (kf + kb):.1f} %)",
        f"# serialised under ncu the two pipelined kernels run at {tf * 136 / fm / 1e6:.0f} / {tf * 176 / bm / 1e6:.0f} GB/s algorithmic; back to back in the bench's timed",
        f"# region the board sits at its power cap ({bench['clocks']}).",
        "# the fallback kernels (scan_*_summary / fold / rerun) appear with ~2 us each: they exit at once when every chunk forgot its start"]
open(os.path.join(P, f"{tag}_launches_summary.txt"), "w").write("\n".join(out) + "\n")
rep = os.path.join(G, "prof_pipe_r2.ncu-rep")
if os.path.exists(rep):
    with open(os.path.join(P, f"{tag}_pipe_kernels_full_raw.csv"), "w") as fh:
        subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=fh, stderr=subprocess.DEVNULL)
print("\n".join(out))
print(json.dumps(traffic["f64"], indent=0)[:900])
print({k: v for k, v in traffic["config5_f64"].items() if k not in ("kernels", "_source")})
for k in traffic["config5_f64"]["kernels"]:
    print("   ", k)
