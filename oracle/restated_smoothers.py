"""CPU oracle, part 2: the task smoothers' parameter setup and output assembly.

TEST INFRASTRUCTURE, NOT PRODUCT CODE (same import rule as ``oracle/restated.py``).
numpy/pandas/scikit-learn restatement of the three L3 drivers of the reference,
array-first (no per-frame DataFrame restacking loops).  Pinned against the
reference's shipped CSV goldens through ``tests/golden`` (see restated.py header).
"""
from __future__ import annotations

import numpy as np
import pandas as pd
from sklearn.decomposition import PCA

from . import restated as core

TRACKER = "ensemble-kalman_tracker"


def dlc_index(keypoint_names):
    """eks/utils.py:4-10."""
    return pd.MultiIndex.from_product(
        [[TRACKER], list(keypoint_names), ["x", "y", "likelihood"]],
        names=["scorer", "bodyparts", "coords"])


# --------------------------------------------------------------------------
# A9  PCA projection + KF parameter setup (multiview_pca_smoother.py:684-731)
# --------------------------------------------------------------------------
def good_frame_mask(evars, quantile):
    """max over observation dims, keep frames <= linear-interp percentile (:685-688)."""
    mx = np.max(evars, axis=1)
    return np.where(mx <= np.percentile(mx, quantile))[0]


def pca_kf_params(fit_rows, good_pcs_of, smooth_param):
    """Shared tail of the multicam / paw setup.

    fit_rows: centred good rows used to fit PCA(3) (:707-708, :189-191).
    good_pcs_of: callable(pca) -> (G,3) PCs of the good frames of *this* track.
    Returns (pca, m0, S0, A, Q, C) following :721-731 / :235-255.
    """
    pca = PCA(n_components=3).fit(fit_rows)
    z = good_pcs_of(pca)
    m0 = np.zeros(3)
    S0 = np.diag(np.var(z, axis=0))                     # ddof=0  (:722)
    A = np.eye(3)
    Q = smooth_param * np.cov((z[1:] - z[:-1]).T)       # consecutive *good* rows, ddof=1 (:726-728)
    C = pca.components_.T                               # (:730)
    return pca, m0, S0, A, Q, C


def multicam_setup(stack, smooth_param, quantile):
    """stack: (M,T,p) with p = 2*n_cams ordered [cam0_x,cam0_y,cam1_x,...].

    Returns dict(y, ev, means, m0, S0, A, Q, C) as at multiview_pca_smoother.py:675-731.
    """
    preds, ev = core.ensemble_arrays(stack, "median")
    good = good_frame_mask(ev, quantile)
    means = preds[good].mean(axis=0)                    # (:698-700)
    y = preds - means
    pca, m0, S0, A, Q, C = pca_kf_params(
        y[good], lambda pc: pc.transform(y)[good], smooth_param)
    return dict(y=y, ev=ev, means=means, m0=m0, S0=S0, A=A, Q=Q, C=C)


def _cam_dfs(ym, means, keypoint, camera_names):
    out = {}
    T = ym.shape[0]
    for c, name in enumerate(camera_names):
        arr = np.column_stack([ym[:, 2 * c] + means[2 * c],
                               ym[:, 2 * c + 1] + means[2 * c + 1],
                               np.full(T, np.nan)])
        out[name + "_df"] = pd.DataFrame(arr, columns=dlc_index([keypoint]))
    return out


def _stack_cameras(markers_list_cameras):
    """(cams)(models) DataFrames -> (M,T,2*cams); first two columns of each frame (:645-659)."""
    n_cam = len(markers_list_cameras)
    M = len(markers_list_cameras[0])
    per = [[markers_list_cameras[c][m].to_numpy()[:, :2] for c in range(n_cam)] for m in range(M)]
    return np.stack([np.concatenate(row, axis=1) for row in per], axis=0).astype(np.float64)


def ensemble_kalman_smoother_multi_cam(markers_list_cameras, keypoint_ensemble, smooth_param,
                                       quantile_keep_pca, camera_names):
    """multiview_pca_smoother.py:611-767."""
    st = multicam_setup(_stack_cameras(markers_list_cameras), smooth_param, quantile_keep_pca)
    R = np.eye(st["C"].shape[0])
    mf, Vf, S = core.filtering_pass(st["y"], st["m0"], st["S0"], st["C"], R, st["A"], st["Q"], st["ev"])
    ms, Vs, _ = core.smooth_backward(st["y"], mf, Vf, S, st["A"], st["Q"], st["C"])
    return _cam_dfs(ms @ st["C"].T, st["means"], keypoint_ensemble, camera_names)


def eks_opti_smoother_multi_cam(markers_list_cameras, keypoint_ensemble, smooth_param,
                                quantile_keep_pca, camera_names, plot=False):
    """multiview_pca_smoother.py:777-933 (plot branch dereferences undefined names; dropped)."""
    st = multicam_setup(_stack_cameras(markers_list_cameras), smooth_param, quantile_keep_pca)
    q = core.kalman_newton_recursive(st["y"], st["m0"], st["S0"], st["A"], st["C"], st["ev"], st["Q"])
    return _cam_dfs(q @ st["C"].T, st["means"], keypoint_ensemble, camera_names)


# --------------------------------------------------------------------------
# paw, two asynchronous views (multiview_pca_smoother.py:34-322)
# --------------------------------------------------------------------------
IMG_WIDTH = 128  # :72


def paw_align(left_np, right_np, ts_left, ts_right):
    """Linear interpolation of the right view onto left timestamps (:70-101).

    left_np (T_l,6), right_np (T_r,6).  Keeps left frames with
    ts_right[0] <= ts <= ts_right[-1] up to the first ts beyond the right range
    (the reference ``break``s there), takes columns [0,1,3,4], flips right x.
    Returns (left (T',4), right (T',4)).
    """
    keep = []
    for i, ts in enumerate(ts_left):
        if ts > ts_right[-1]:
            break
        if ts < ts_right[0]:
            continue
        keep.append(i)
    keep = np.asarray(keep, dtype=np.int64)
    cols = [0, 1, 3, 4]
    left = left_np[keep][:, cols].astype(np.float64)
    t = ts_left[keep]
    # scipy interp1d(kind='linear') == piecewise-linear between bracketing samples
    hi = np.clip(np.searchsorted(ts_right, t, side="left"), 1, len(ts_right) - 1)
    lo = hi - 1
    slope = (right_np[hi][:, cols] - right_np[lo][:, cols]) / (ts_right[hi] - ts_right[lo])[:, None]
    right = slope * (t - ts_right[lo])[:, None] + right_np[lo][:, cols]
    right[:, 0] = IMG_WIDTH - right[:, 0]
    right[:, 2] = IMG_WIDTH - right[:, 2]
    return left, right


def paw_setup(left_stack, right_stack, smooth_param, quantile):
    """left/right_stack: (M,T',4) aligned [paw_l_x,paw_l_y,paw_r_x,paw_r_y] per view (:115-255)."""
    lp, lv = core.ensemble_arrays(left_stack, "median")
    rp, rv = core.ensemble_arrays(right_stack, "median")
    good = good_frame_mask(np.hstack([lv, rv]), quantile)
    y_paw = [np.hstack([lp[:, :2], rp[:, :2]]), np.hstack([lp[:, 2:4], rp[:, 2:4]])]
    ev_paw = [np.hstack([lv[:, :2], rv[:, :2]]), np.hstack([lv[:, 2:4], rv[:, 2:4]])]
    G = good.shape[0]
    stacked = np.empty((2 * G, 4))                      # interleaved paw rows (:135-151)
    stacked[0::2] = y_paw[0][good]
    stacked[1::2] = y_paw[1][good]
    means = stacked.mean(axis=0)                        # (:173-177)
    tracks = []
    for k in range(2):
        y = y_paw[k] - means
        pca, m0, S0, A, Q, C = pca_kf_params(
            stacked - means, lambda pc, y=y: pc.transform(y)[good], smooth_param)
        tracks.append(dict(y=y, ev=ev_paw[k], m0=m0, S0=S0, A=A, Q=Q, C=C))
    return tracks, means


def ensemble_kalman_smoother_paw_asynchronous(markers_list_left_cam, markers_list_right_cam,
                                              timestamps_left_cam, timestamps_right_cam,
                                              keypoint_names, smooth_param, quantile_keep_pca, newton=False):
    """multiview_pca_smoother.py:34-322; newton=True is its opti twin (:326-608, Newton filter at :552)."""
    aligned = [paw_align(l.to_numpy(), r.to_numpy(), timestamps_left_cam, timestamps_right_cam)
               for l, r in zip(markers_list_left_cam, markers_list_right_cam)]
    left_stack = np.stack([a[0] for a in aligned])
    right_stack = np.stack([a[1] for a in aligned])
    tracks, means = paw_setup(left_stack, right_stack, smooth_param, quantile_keep_pca)
    ym = []
    for st in tracks:
        R = np.eye(4)
        if newton:
            ms = core.kalman_newton_recursive(st["y"], st["m0"], st["S0"], st["A"], st["C"], st["ev"], st["Q"])
        else:
            mf, Vf, S = core.filtering_pass(st["y"], st["m0"], st["S0"], st["C"], R, st["A"], st["Q"], st["ev"])
            ms, _, _ = core.smooth_backward(st["y"], mf, Vf, S, st["A"], st["Q"], st["C"])
        ym.append(ms @ st["C"].T + means)
    T = ym[0].shape[0]
    nan = np.full(T, np.nan)
    # left view: paw_l then paw_r; right view: swapped and un-flipped (:296-320)
    left = np.column_stack([ym[0][:, 0], ym[0][:, 1], nan, ym[1][:, 0], ym[1][:, 1], nan])
    right = np.column_stack([IMG_WIDTH - ym[1][:, 2], ym[1][:, 3], nan,
                             IMG_WIDTH - ym[0][:, 2], ym[0][:, 3], nan])
    idx = dlc_index(keypoint_names)
    return {"left_df": pd.DataFrame(left, columns=idx), "right_df": pd.DataFrame(right, columns=idx)}


def eks_opti_smoother_paw_asynchronous(markers_list_left_cam, markers_list_right_cam, timestamps_left_cam,
                                       timestamps_right_cam, keypoint_names, smooth_param, quantile_keep_pca):
    return ensemble_kalman_smoother_paw_asynchronous(markers_list_left_cam, markers_list_right_cam, timestamps_left_cam,
                                                     timestamps_right_cam, keypoint_names, smooth_param,
                                                     quantile_keep_pca, newton=True)


# --------------------------------------------------------------------------
# pupil (pupil_smoother.py:14-79 geometry, :82-223 smoother, :227-364 opti twin)
# --------------------------------------------------------------------------
PUPIL_KEYS = ['pupil_top_r_x', 'pupil_top_r_y', 'pupil_bottom_r_x', 'pupil_bottom_r_y',
              'pupil_right_r_x', 'pupil_right_r_y', 'pupil_left_r_x', 'pupil_left_r_y']   # :102-103
PUPIL_C = np.asarray([[0, 1, 0], [-.5, 0, 1], [0, 1, 0], [.5, 0, 1],
                      [.5, 1, 0], [0, 0, 1], [-.5, 1, 0], [0, 0, 1]], dtype=np.float64)   # :150-153


def pupil_location(P):
    """P: (T,8) in PUPIL_KEYS order -> (T,2) centre (pupil_smoother.py:14-41)."""
    t, b, r, l = P[:, 0:2], P[:, 2:4], P[:, 4:6], P[:, 6:8]
    x1 = np.nanmedian(np.column_stack([t[:, 0], b[:, 0]]), axis=1)
    x2 = np.median(np.column_stack([r[:, 0], l[:, 0]]), axis=1)
    y1 = np.median(np.column_stack([t[:, 1], b[:, 1]]), axis=1)
    y2 = np.nanmedian(np.column_stack([r[:, 1], l[:, 1]]), axis=1)
    return np.column_stack([np.nanmedian(np.column_stack([x1, x2]), axis=1),
                            np.nanmedian(np.column_stack([y1, y2]), axis=1)])


def pupil_diameter(P):
    """nanmedian of the six diameter estimates (pupil_smoother.py:44-68)."""
    t, b, r, l = P[:, 0:2], P[:, 2:4], P[:, 4:6], P[:, 6:8]
    d = [np.linalg.norm(t - b, axis=1), np.linalg.norm(l - r, axis=1)]
    for u, v in [(t, l), (t, r), (b, l), (b, r)]:
        d.append(np.linalg.norm(u - v, axis=1) * 2 ** 0.5)
    return np.nanmedian(np.asarray(d), axis=0)


def pupil_setup(stack, A):
    """stack (M,T,8) in PUPIL_KEYS order; A (3,3) diagonal transition (:104-172)."""
    preds, ev = core.ensemble_arrays(stack, "median")
    loc = pupil_location(preds)
    diam = pupil_diameter(preds)
    mx, my = np.mean(loc[:, 0]), np.mean(loc[:, 1])
    xo, yo = loc[:, 0] - mx, loc[:, 1] - my
    m0 = np.asarray([np.mean(diam), 0.0, 0.0])
    v = np.asarray([np.var(diam), np.var(xo), np.var(yo)])
    S0 = np.diag(v)
    Q = np.diag(v * (1.0 - np.diag(A) ** 2))
    y = preds.copy()
    y[:, 0::2] -= mx
    y[:, 1::2] -= my
    return dict(y=y, ev=ev, m0=m0, S0=S0, A=np.asarray(A, dtype=np.float64), Q=Q, C=PUPIL_C.copy(),
                mean_x=mx, mean_y=my)


def _pupil_frames(ym, lat, st, keypoint_names, tracker_name):
    T = ym.shape[0]
    full = ym.copy()
    full[:, 0::2] += st["mean_x"]          # keys containing 'x' (add_mean_to_array :71-79;
    full[:, 1::2] += st["mean_y"]          # every PUPIL_KEY contains '_x' or '_y', none both)
    nan = np.full(T, np.nan)
    order = [0, 4, 2, 6]                   # top, right, bottom, left (:199-202)
    cols = []
    for o in order:
        cols += [full[:, o], full[:, o + 1], nan]
    markers_df = pd.DataFrame(np.column_stack(cols), columns=dlc_index(keypoint_names))
    latents = np.column_stack([lat[:, 0], lat[:, 1] + st["mean_x"], lat[:, 2] + st["mean_y"]])
    idx = pd.MultiIndex.from_arrays([[tracker_name] * 3, ['diameter', 'com_x', 'com_y']],
                                    names=('scorer', 'latent'))
    return {"markers_df": markers_df, "latents_df": pd.DataFrame(latents, columns=idx)}


def _pupil_stack(markers_list):
    return np.stack([np.column_stack([np.asarray(df[k], dtype=np.float64) for k in PUPIL_KEYS])
                     for df in markers_list])


def ensemble_kalman_smoother_pupil(markers_list, keypoint_names, tracker_name, state_transition_matrix):
    st = pupil_setup(_pupil_stack(markers_list), state_transition_matrix)
    R = np.eye(8)
    mf, Vf, S = core.filtering_pass(st["y"], st["m0"], st["S0"], st["C"], R, st["A"], st["Q"], st["ev"])
    ms, _, _ = core.smooth_backward(st["y"], mf, Vf, S, st["A"], st["Q"], st["C"])
    return _pupil_frames(ms @ st["C"].T, ms, st, keypoint_names, tracker_name)


def eks_opti_pupil_latents(markers_list):
    """Latents of eks_opti_smoother_pupil (pupil_smoother.py:227-364): hard-coded A=0.99 I (:301-305).

    Only latents_df is meaningful there (markers_df is rebuilt from a CSV of the standard
    smoother, Appendix C.10), so the oracle returns the (T,3) latents with means added back.
    """
    st = pupil_setup(_pupil_stack(markers_list), np.diag([0.99, 0.99, 0.99]))
    q = core.kalman_newton_recursive(st["y"], st["m0"], st["S0"], st["A"], st["C"], st["ev"], st["Q"])
    return np.column_stack([q[:, 0], q[:, 1] + st["mean_x"], q[:, 2] + st["mean_y"]])


def eks_opti_smoother_pupil(markers_list, keypoint_names, tracker_name, state_transition_matrix=None):
    """eks_opti_smoother_pupil with its default plot=True (pupil_smoother.py:227-364).  latents_df: the Newton
    latents.  markers_df: the reference overwrites q with the latents CSV of the STANDARD smoother (:325-329; the
    shipped kalman_smoothed_latents.csv = the standard smoother at A = 0.99 I, means already added to com_x/com_y)
    and forms B q + means from it (:332-337), so the means enter twice.  Reproduced as shipped."""
    A = np.diag([0.99, 0.99, 0.99])
    st = pupil_setup(_pupil_stack(markers_list), A)
    q = core.kalman_newton_recursive(st["y"], st["m0"], st["S0"], st["A"], st["C"], st["ev"], st["Q"])
    std = ensemble_kalman_smoother_pupil(markers_list, keypoint_names, tracker_name, A)["latents_df"].to_numpy()
    return _pupil_frames(std @ st["C"].T, q, st, keypoint_names, tracker_name)
