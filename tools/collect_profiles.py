"""Copy the evidence of the last tools/gpu_round.sh run from gpurun_out/ (scratch) into profiles/ (tracked):
bench lines, ncu launch list + summary, full-size DRAM-bytes capture -> traffic.json, ncu --set full raw page."""
import collections, csv, json, os, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
for src, dst in (("bench_f64.json", f"{tag}_bench_f64.json"), ("bench_f32.json", f"{tag}_bench_f32.json"),
                 ("bench_ref.json", f"{tag}_bench_reference.json"), ("dram_full.csv", f"{tag}_dram_bytes_full_size.csv"),
                 ("launches.csv", f"{tag}_launches.csv"), ("scan_time.log", f"{tag}_scan_time.txt"),
                 ("setup_time.log", f"{tag}_setup_time.txt")):
    if os.path.exists(os.path.join(G, src)):
        shutil.copy(os.path.join(G, src), os.path.join(P, dst))


def table(path):
    rows = list(csv.reader(open(path)))
    i = [k for k, r in enumerate(rows) if r and r[0] == "ID"][0]
    return rows[i], rows[i + 1:]


h, rows = table(os.path.join(G, "dram_full.csv"))
kn, mn, mv = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value")
acc = collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows:
    if len(r) > mv:
        acc["fwd" if "kf_forward" in r[kn] else "bwd"][r[mn]].append(float(r[mv].replace(",", "")))
mean = lambda k, m: sum(acc[k][m]) / len(acc[k][m])
bench = json.loads(open(os.path.join(G, "bench_f64.json")).read().strip().splitlines()[-1])
tf = bench["config"]["tracks_per_gpu"] * bench["config"]["frames"]
traffic = {"_source": f"profiles/{tag}_dram_bytes_full_size.csv: ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,"
                      "gpu__time_duration.sum on `python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e` (f64, "
                      f"{bench['config']['tracks_per_gpu']} tracks x {bench['config']['frames']} frames), mean per launch",
           "f64": {"kernel": "rts_backward_pipe_kernel", "dram_bytes": int(mean("bwd", "dram__bytes_read.sum") + mean("bwd", "dram__bytes_write.sum")),
                   "read": int(mean("bwd", "dram__bytes_read.sum")), "write": int(mean("bwd", "dram__bytes_write.sum")),
                   "algorithmic": tf * 176, "ncu_duration_ms": mean("bwd", "gpu__time_duration.sum") / 1e6,
                   "kf_forward_pipe_kernel": {"dram_bytes": int(mean("fwd", "dram__bytes_read.sum") + mean("fwd", "dram__bytes_write.sum")),
                                              "read": int(mean("fwd", "dram__bytes_read.sum")), "write": int(mean("fwd", "dram__bytes_write.sum")),
                                              "algorithmic": tf * 136, "ncu_duration_ms": mean("fwd", "gpu__time_duration.sum") / 1e6}}}
json.dump(traffic, open(os.path.join(P, "traffic.json"), "w"), indent=1)

h, rows = table(os.path.join(G, "launches.csv"))
kn, mv = h.index("Kernel Name"), h.index("Metric Value")
agg = collections.OrderedDict()
for r in rows:
    if len(r) > mv:
        agg.setdefault(r[kn].split("(")[0][:100], []).append(float(r[mv].replace(",", "")))
out = ["# ncu --metrics gpu__time_duration.sum --clock-control none -k regex:kernel python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e",
       "# (B200, f64, 8192 tracks x 100000 frames; serialised cold-cache times: compare SHARES)", "# launches  total_ms  ms/launch  kernel"]
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    out.append(f"{len(v):6d} {sum(v) / 1e6:10.2f} {sum(v) / len(v) / 1e6:9.3f}  {k}")
f = [v for k, v in agg.items() if "kf_forward_pipe" in k][0]; b = [v for k, v in agg.items() if "rts_backward_pipe" in k][0]
fm, bm = sum(f) / len(f) / 1e6, sum(b) / len(b) / 1e6
kf, kb = bench["roofline"]["kernels"]["kf_forward"]["ms"], bench["roofline"]["kernels"]["rts_backward"]["ms"]
out += [f"# timed step = 1 x kf_forward_pipe_kernel + 1 x rts_backward_pipe_kernel: shares {100 * fm / (fm + bm):.1f}% / {100 * bm / (fm + bm):.1f}% "
        f"(bench CUDA events, same box: {kf:.2f} / {kb:.2f} ms = {100 * kf / (kf + kb):.1f}% / {100 * kb / (kf + kb):.1f}%)",
        f"# serialised under ncu the two kernels run at {tf * 136 / fm / 1e6:.0f} / {tf * 176 / bm / 1e6:.0f} GB/s algorithmic; back to back in the bench's timed",
        f"# region the board sits at its power cap ({bench['clocks']}).",
        "# at::tensor_kernel_scan_outer_dim / distribution_* / elementwise = torch synthesising the input tensor (outside the timed region)"]
open(os.path.join(P, f"{tag}_launches_summary.txt"), "w").write("\n".join(out) + "\n")
rep = os.path.join(G, "prof_pipe.ncu-rep")
if os.path.exists(rep):
    with open(os.path.join(P, f"{tag}_pipe_kernels_full_raw.csv"), "w") as fh:
        subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], stdout=fh, stderr=subprocess.DEVNULL)
print("\n".join(out[-5:])); print(json.dumps(traffic["f64"])[:400])
for n in ("bench_f64", "bench_f32", "bench_ref"):
    d = json.loads(open(os.path.join(G, n + ".json")).read().strip().splitlines()[-1])
    r = d.get("roofline", {})
    print(n, "%.4g" % d["value"], r.get("step_frac"), r.get("frac"), {k: (round(v["ms"], 2), round(v["frac"], 3)) for k, v in r.get("kernels", {}).items()},
          d.get("e2e", {}).get("value"), d.get("cpu_baseline", {}) and d["cpu_baseline"]["value"], d.get("clocks"), r.get("sustained_copy"))
