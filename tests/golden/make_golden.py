"""Generate tests/golden/*.npz by running the UNMODIFIED reference (erialc-cal/eks).

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
The reference is pure Python and cannot travel to the GPU box, so its inputs and
outputs are frozen here as small fixtures.  Each fixture stores (a) the raw inputs
as plain arrays, (b) what the reference functions return on them, and (c) where the
reference ships its own output CSV, those shipped values too (so a test can show
reference-run == shipped golden == oracle == CUDA path).
"""
import os
import sys
import types

import numpy as np
import pandas as pd

REF = os.environ.get("EKS_REFERENCE", "/root/reference")
OUT = os.path.dirname(os.path.abspath(__file__))

# matplotlib / seaborn are absent here; the reference imports them at module scope
# (eks/newton_eks.py:7-8).  Empty stubs are enough.
for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.gridspec", "seaborn"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.path.insert(0, REF)

import eks.ensemble_kalman as ek                      # noqa: E402
import eks.multiview_pca_smoother as mv               # noqa: E402
import eks.pupil_smoother as ps                       # noqa: E402
from eks.newton_eks import kalman_newton_recursive    # noqa: E402
from eks.utils import convert_lp_dlc                  # noqa: E402


def load_dir(csv_dir, pred=lambda f: True):
    out = []
    for f in sorted(os.listdir(csv_dir)):
        if not f.endswith("csv") or not pred(f):
            continue
        df = pd.read_csv(os.path.join(csv_dir, f), header=[0, 1, 2], index_col=0)
        kp = [c[1] for c in df.columns[::3]]
        out.append((convert_lp_dlc(df, kp, model_name=df.columns[0][0]), kp))
    return [o[0] for o in out], out[0][1]


def by_camera(markers_list, keypoint, camera_names):
    cams = [[] for _ in camera_names]
    for df in markers_list:
        for c, cam in enumerate(camera_names):
            keys = [k for k in df.keys() if cam in k and "likelihood" not in k and keypoint in k]
            cams[c].append(df[keys])
    return cams


def stack_of(cams):
    M = len(cams[0])
    return np.stack([np.concatenate([cams[c][m].to_numpy()[:, :2] for c in range(len(cams))], 1)
                     for m in range(M)]).astype(np.float64)


def core_case(y, m0, S0, C, A, Q, ev):
    R = np.eye(C.shape[0])
    mf, Vf, S = ek.filtering_pass(y, m0, S0, C, R, A, Q, ev)
    ms, Vs, CV = ek.smooth_backward(y, mf, Vf, S, A, Q, C)
    return dict(y=y, ev=ev, m0=m0, S0=S0, C=C, A=A, Q=Q, R_after=R, mf=mf, Vf=Vf, S=S, ms=ms, Vs=Vs, CV=CV)


def mirror_mouse():
    d = os.path.join(REF, "data/mirror-mouse")
    ml, _ = load_dir(d)
    cams = ["top", "bot"]
    paws = ["paw1LH", "paw2LF", "paw3RF", "paw4RH"]
    shipped = pd.read_csv(os.path.join(d, "output/eks.csv"), header=[0, 1, 2], index_col=0)
    shipped_o = pd.read_csv(os.path.join(d, "output/eks_opti.csv"), header=[0, 1, 2], index_col=0)
    z = {}
    for kp in paws:
        bc = by_camera(ml, kp, cams)
        z[f"{kp}/stack"] = stack_of(bc)
        std = mv.ensemble_kalman_smoother_multi_cam(bc, kp, 0.01, 25, cams)
        opt = mv.eks_opti_smoother_multi_cam(bc, kp, 0.01, 25, cams, plot=False)
        pick = lambda dd: np.column_stack([dd[f"{c}_df"].to_numpy()[:, :2] for c in cams])
        z[f"{kp}/ref_std"] = pick(std)
        z[f"{kp}/ref_opti"] = pick(opt)
        tr = "ensemble-kalman_tracker"
        z[f"{kp}/shipped_std"] = np.column_stack([shipped[(tr, f"{kp}_{c}", xy)] for c in cams for xy in "xy"])
        z[f"{kp}/shipped_opti"] = np.column_stack([shipped_o[(tr, f"{kp}_{c}", xy)] for c in cams for xy in "xy"])
    np.savez_compressed(os.path.join(OUT, "mirror_mouse.npz"), **z)
    print("mirror_mouse", {k: float(np.abs(z[f'{k}/ref_std'] - z[f'{k}/shipped_std']).max()) for k in paws},
          {k: float(np.abs(z[f'{k}/ref_opti'] - z[f'{k}/shipped_opti']).max()) for k in paws})


def pupil():
    d = os.path.join(REF, "data/ibl-pupil")
    ml, kp = load_dir(d)
    keys = ['pupil_top_r_x', 'pupil_top_r_y', 'pupil_bottom_r_x', 'pupil_bottom_r_y',
            'pupil_right_r_x', 'pupil_right_r_y', 'pupil_left_r_x', 'pupil_left_r_y']
    z = {"stack": np.stack([np.column_stack([df[k] for k in keys]) for df in ml]).astype(np.float64)}
    z["keypoint_names"] = np.asarray(kp)
    for tag, A in (("a99", np.diag([.99, .99, .99])), ("adef", np.diag([.9999, .999, .999]))):
        out = ps.ensemble_kalman_smoother_pupil(ml, kp, "ensemble-kalman_tracker", A)
        z[f"{tag}/markers"] = out["markers_df"].to_numpy()
        z[f"{tag}/latents"] = out["latents_df"].to_numpy()
    t = os.path.join(REF, "data/misc/pupil-test")
    z["shipped/latents"] = pd.read_csv(os.path.join(t, "kalman_smoothed_latents.csv"), header=[0, 1], index_col=0).to_numpy()
    z["shipped/markers"] = pd.read_csv(os.path.join(t, "kalman_smoothed_pupil_traces.csv"), header=[0, 1, 2], index_col=0).to_numpy()
    z["shipped/opti_latents"] = pd.read_csv(os.path.join(t, "opti_eks_latents.csv"), header=[0, 1], index_col=0).to_numpy()
    # the reference's opti pupil markers (pupil_smoother.py:325-346): built from the STANDARD smoother's latents CSV
    z["shipped/opti_markers"] = pd.read_csv(os.path.join(t, "opti_eks_pupil_traces.csv"), header=[0, 1, 2], index_col=0).to_numpy()
    np.savez_compressed(os.path.join(OUT, "pupil.npz"), **z)
    print("pupil", np.nanmax(np.abs(z["a99/latents"] - z["shipped/latents"])),
          np.nanmax(np.abs(z["a99/markers"] - z["shipped/markers"])))


def paw():
    d = os.path.join(REF, "data/ibl-paw")
    left, kp = load_dir(d, lambda f: "left" in f)
    right, _ = load_dir(d, lambda f: "right" in f)
    cols = {'paw_l_x': 'paw_r_x', 'paw_l_y': 'paw_r_y', 'paw_l_likelihood': 'paw_r_likelihood',
            'paw_r_x': 'paw_l_x', 'paw_r_y': 'paw_l_y', 'paw_r_likelihood': 'paw_l_likelihood'}
    right = [df.rename(columns=cols).loc[:, cols.keys()] for df in right]   # scripts/multiview_paw_example.py:83-92
    tl = np.load([os.path.join(d, f) for f in os.listdir(d) if "timestamps.left" in f][0], allow_pickle=True)
    tr = np.load([os.path.join(d, f) for f in os.listdir(d) if "timestamps.right" in f][0], allow_pickle=True)
    out = mv.ensemble_kalman_smoother_paw_asynchronous(left, right, tl, tr, kp, 0.01, 25)
    z = dict(left=np.stack([df.to_numpy() for df in left]).astype(np.float64),
             right=np.stack([df.to_numpy() for df in right]).astype(np.float64),
             ts_left=np.asarray(tl, dtype=np.float64), ts_right=np.asarray(tr, dtype=np.float64),
             keypoint_names=np.asarray(kp),
             left_df=out["left_df"].to_numpy(), right_df=out["right_df"].to_numpy())
    opti = mv.eks_opti_smoother_paw_asynchronous(left, right, tl, tr, kp, 0.01, 25)      # :326, Newton filter at :552
    z["opti_left_df"] = opti["left_df"].to_numpy()
    z["opti_right_df"] = opti["right_df"].to_numpy()
    np.savez_compressed(os.path.join(OUT, "paw.npz"), **z)
    print("paw", z["left_df"].shape, z["right_df"].shape, z["opti_left_df"].shape)


def fish(n_clips=10):
    base = os.path.join(REF, "data/misc/mirror-fish_ensemble-predictions")
    cams = ["main", "top", "right"]
    z = {}
    sessions = sorted(os.listdir(os.path.join(base, "network_0")))
    clips = []
    for s in sessions:
        sd = os.path.join(base, "network_0", s)
        if os.path.isdir(sd):
            clips += [(s, f) for f in sorted(os.listdir(sd)) if f.endswith(".csv")]
    step = max(1, len(clips) // n_clips)
    worst = 0.0
    for (s, f) in clips[::step][:n_clips]:
        ml = []
        for k in range(5):
            df = pd.read_csv(os.path.join(base, f"network_{k}", s, f), header=[0, 1, 2], index_col=0)
            kps = [c[1] for c in df.columns[::3]]
            ml.append(convert_lp_dlc(df, kps, model_name=df.columns[0][0]))
        shipped = pd.read_csv(os.path.join(base, "eks", s, f), header=[0, 1, 2], index_col=0)
        bodyparts = sorted({k.rsplit("_", 2)[0] for k in ml[0].keys() if k.endswith("_main_x")})
        for bp in bodyparts:
            bc = by_camera(ml, bp, cams)
            if any(len(c[0].columns) != 2 for c in bc):
                continue
            out = mv.ensemble_kalman_smoother_multi_cam(bc, bp, 0.01, 50, cams)
            key = f"{s}|{f}|{bp}"
            z[key + "/stack"] = stack_of(bc)
            z[key + "/ref"] = np.column_stack([out[f"{c}_df"].to_numpy()[:, :2] for c in cams])
            tr = shipped.columns[0][0]          # test_script.py writes under the input tracker name
            try:
                z[key + "/shipped"] = np.column_stack([shipped[(tr, f"{bp}_{c}", xy)] for c in cams for xy in "xy"])
                worst = max(worst, float(np.abs(z[key + "/ref"] - z[key + "/shipped"]).max()))
            except KeyError:
                pass
    np.savez_compressed(os.path.join(OUT, "fish.npz"), **z)
    print("fish tracks", len(z) // 3, "worst ref-vs-shipped", worst)


def core_kats():
    rng = np.random.default_rng(20240607)
    z = {}
    for name, (T, p, n) in {"p4": (257, 4, 3), "p6": (100, 6, 3), "p8": (129, 8, 3), "T1": (1, 4, 3), "T2": (2, 4, 3),
                            "n2p4": (64, 4, 2), "n4p6": (64, 6, 4), "n6p8": (64, 8, 6)}.items():
        C = np.linalg.qr(rng.normal(size=(p, n)))[0]
        A = np.diag(rng.uniform(0.9, 1.0, size=n))
        L = rng.normal(size=(n, n)) * 0.1
        Q = L @ L.T + 0.01 * np.eye(n)
        S0 = np.diag(rng.uniform(0.5, 2.0, size=n))
        m0 = rng.normal(size=n)
        x = np.cumsum(rng.normal(size=(T, n)) * 0.1, axis=0)
        sd = np.exp(rng.normal(np.log(0.5), 0.8, size=(T, p)))
        y = x @ C.T + sd * rng.normal(size=(T, p))
        ev = sd ** 2 / 5
        if T < 2:
            R = np.eye(p)
            mf, Vf, S = ek.filtering_pass(y, m0, S0, C, R, A, Q, ev)
            case = dict(y=y, ev=ev, m0=m0, S0=S0, C=C, A=A, Q=Q, R_after=R, mf=mf, Vf=Vf, S=S)
        else:
            case = core_case(y, m0, S0, C, A, Q, ev)
            case["newton"] = kalman_newton_recursive(y, m0, S0, A, C, ev, Q, max_iter=1)
        # general (non-diagonal) A and a dense R with off-diagonals (signature allows both)
        for k, v in case.items():
            z[f"{name}/{k}"] = v
    # dense-A / dense-R case
    T, p, n = 80, 4, 3
    C = rng.normal(size=(p, n)); A = np.eye(n) * 0.95 + 0.03 * rng.normal(size=(n, n))
    Q = 0.05 * np.eye(n); S0 = np.eye(n); m0 = np.zeros(n)
    y = rng.normal(size=(T, p)); ev = np.exp(rng.normal(size=(T, p)))
    R = np.eye(p) + 0.1; R = 0.5 * (R + R.T)
    R_in = R.copy()
    mf, Vf, S = ek.filtering_pass(y, m0, S0, C, R, A, Q, ev)
    ms, Vs, CV = ek.smooth_backward(y, mf, Vf, S, A, Q, C)
    for k, v in dict(y=y, ev=ev, m0=m0, S0=S0, C=C, A=A, Q=Q, R_in=R_in, R_after=R, mf=mf, Vf=Vf, S=S, ms=ms, Vs=Vs, CV=CV).items():
        z[f"denseAR/{k}"] = v
    # ensemble KATs via the reference's DataFrame API: odd/even M, NaN, mean mode
    for name, (M, T, K) in {"m5": (5, 64, 6), "m4": (4, 33, 3), "m1": (1, 8, 2), "m8": (8, 40, 4), "m3nan": (3, 16, 2)}.items():
        st = rng.normal(size=(M, T, K)) * 10
        if "nan" in name:
            st[1, 3, 0] = np.nan
            st[:, 5, 1] = np.nan
        keys = [f"k{i}" for i in range(K)]
        dfs = [pd.DataFrame(st[m], columns=keys) for m in range(M)]
        for mode in ("median", "mean"):
            pr, va, stk, *_ = ek.ensemble(dfs, keys, mode)
            z[f"ens_{name}/{mode}_preds"] = pr
            z[f"ens_{name}/{mode}_vars"] = va
        z[f"ens_{name}/stack"] = st
        z[f"ens_{name}/stacks_out"] = np.ascontiguousarray(stk)
    np.savez_compressed(os.path.join(OUT, "core_kats.npz"), **z)
    print("core_kats", len(z))


def extra_kats():
    """Round-2 additions (own file and own seed so the round-1 fixtures stay byte-stable): a zero ensemble variance,
    shapes outside the template instantiations, kalman_newton_recursive(max_iter>1), smooth_backward with a
    caller-modified S, a 20-model ensemble, and a five-camera / ten-model multi-camera smoother call."""
    rng = np.random.default_rng(777)
    z = {}

    def synth(T, p, n):
        C = np.linalg.qr(rng.normal(size=(p, n)))[0] if p >= n else rng.normal(size=(p, n))
        A = np.diag(rng.uniform(0.9, 1.0, size=n))
        L = rng.normal(size=(n, n)) * 0.1
        Q = L @ L.T + 0.01 * np.eye(n)
        S0 = np.diag(rng.uniform(0.5, 2.0, size=n))
        m0 = rng.normal(size=n)
        x = np.cumsum(rng.normal(size=(T, n)) * 0.1, axis=0)
        sd = np.exp(rng.normal(np.log(0.5), 0.8, size=(T, p)))
        return x @ C.T + sd * rng.normal(size=(T, p)), m0, S0, C, A, Q, sd ** 2 / 5

    y, m0, S0, C, A, Q, ev = synth(50, 4, 3)
    ev[10, 2] = 0.0; ev[20, 0] = 0.0; ev[21, 0] = 0.0             # every model agreed to the last bit
    case = core_case(y, m0, S0, C, A, Q, ev)
    # (no Newton twin here: kalman_newton_recursive inverts D_t itself, newton_eks.py:136, so a zero variance is inf)
    for k, v in case.items():
        z[f"zerovar/{k}"] = v
    for name, (T, p, n) in {"n5p10": (40, 10, 5), "n5p4": (40, 4, 5), "n7p12": (30, 12, 7), "n3p10": (60, 10, 3)}.items():
        y, m0, S0, C, A, Q, ev = synth(T, p, n)
        case = core_case(y, m0, S0, C, A, Q, ev)
        case["newton"] = kalman_newton_recursive(y, m0, S0, A, C, ev, Q, max_iter=1)
        for k, v in case.items():
            z[f"{name}/{k}"] = v
    y, m0, S0, C, A, Q, ev = synth(120, 4, 3)
    q3, loss3 = kalman_newton_recursive(y, m0, S0, A, C, ev, Q, max_iter=3)
    for k, v in dict(y=y, m0=m0, S0=S0, C=C, A=A, Q=Q, ev=ev, q3=q3, loss3=loss3,
                     q1=kalman_newton_recursive(y, m0, S0, A, C, ev, Q, max_iter=1)).items():
        z[f"newton3/{k}"] = v
    case = core_case(y, m0, S0, C, A, Q, ev)
    S_own = case["S"] * 1.1 + 0.01 * np.eye(3)                    # the reference reads S, never Q (:158-160)
    ms, Vs, CV = ek.smooth_backward(y, case["mf"], case["Vf"], S_own, A, Q, C)
    for k, v in dict(y=y, mf=case["mf"], Vf=case["Vf"], S_own=S_own, A=A, Q=Q, C=C, ms=ms, Vs=Vs, CV=CV).items():
        z[f"ownS/{k}"] = v
    st = rng.normal(size=(20, 30, 3)) * 10
    keys = ["k0", "k1", "k2"]
    dfs = [pd.DataFrame(st[m], columns=keys) for m in range(20)]
    for mode in ("median", "mean"):
        pr, va, *_ = ek.ensemble(dfs, keys, mode)
        z[f"ens_m20/{mode}_preds"] = pr
        z[f"ens_m20/{mode}_vars"] = va
    z["ens_m20/stack"] = st
    # five cameras (p = 10), ten ensemble members, through the reference's multi-camera smoother
    M, T, ncam = 10, 300, 5
    Cx = rng.normal(size=(2 * ncam, 3))
    x = np.cumsum(rng.normal(size=(T, 3)) * 0.3, axis=0)
    clean = x @ Cx.T + rng.normal(size=2 * ncam) * 50
    sd = np.exp(rng.normal(np.log(0.5), 0.8, size=(T, 2 * ncam)))
    stack = clean[None] + sd[None] * rng.normal(size=(M, T, 2 * ncam))
    cams = [f"cam{c}" for c in range(ncam)]
    mlc = [[pd.DataFrame(np.column_stack([stack[m][:, 2 * c], stack[m][:, 2 * c + 1], np.ones(T)]),
                         columns=[f"kp_{cams[c]}_x", f"kp_{cams[c]}_y", f"kp_{cams[c]}_likelihood"]) for m in range(M)]
           for c in range(ncam)]
    out = mv.ensemble_kalman_smoother_multi_cam(mlc, "kp", 0.01, 25, cams)
    z["multicam5/stack"] = stack
    z["multicam5/ref"] = np.column_stack([out[f"{c}_df"].to_numpy()[:, :2] for c in cams])
    np.savez_compressed(os.path.join(OUT, "extra_kats.npz"), **z)
    print("extra_kats", len(z))


if __name__ == "__main__":
    if sys.argv[1:] == ["extra"]:
        extra_kats(); pupil(); paw()
        sys.exit(0)
    if sys.argv[1:] == ["fish"]:
        fish()
        sys.exit(0)
    core_kats()
    extra_kats()
    mirror_mouse()
    pupil()
    paw()
    fish()
