"""Latency of the three real-data configs (BASELINE configs[0..2]) on the golden inputs: the drop-in smoothers on
the GPU next to the oracle (numpy restatement of the reference) on this box's host.  Correctness of the same
calls is covered by tests/test_gpu_smoothers.py; sizes are far too small to load a GPU (B*T <= 8000)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, pandas as pd, torch
from conftest import load_golden, cases_of
from oracle import restated_smoothers as osm
from eks_b200.multiview_pca_smoother import smooth_session_multi_cam


def best(fn, n=5):
    ts = []
    for _ in range(n):
        t = time.perf_counter(); fn(); torch.cuda.synchronize(); ts.append(time.perf_counter() - t)
    return min(ts) * 1e3


z = cases_of(load_golden("mirror_mouse.npz")); kps = sorted(z)
stack = np.stack([z[k]["stack"] for k in kps], axis=2)
smooth_session_multi_cam(stack, 0.01, 25)
gpu = best(lambda: smooth_session_multi_cam(stack, 0.01, 25))
def cpu_mm():
    for k in kps:
        s = osm.multicam_setup(z[k]["stack"], 0.01, 25)
        from oracle import restated as core
        mf, Vf, S = core.filtering_pass(s["y"], s["m0"], s["S0"], s["C"], np.eye(4), s["A"], s["Q"], s["ev"])
        core.smooth_backward(s["y"], mf, Vf, S, s["A"], s["Q"], s["C"])
cpu = best(cpu_mm, 2)
print(f"mirror-mouse  4 keypoints x 2000 frames x 2 cams: GPU session call {gpu:8.2f} ms | CPU oracle {cpu:9.1f} ms | x{cpu/gpu:.0f}")
