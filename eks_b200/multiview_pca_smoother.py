"""Drop-in for ``eks/multiview_pca_smoother.py`` (erialc-cal/eks): the multi-camera and the
asynchronous two-view paw smoothers and their Newton ("opti") twins, same names, arguments
and DataFrame outputs.

Host code here only restacks inputs, fits the 3-component PCA that defines (C, S0, Q)
(multiview_pca_smoother.py:684-731) and assembles DataFrames.  The ensemble reduction, the
Kalman filter, the RTS smoother and the Newton recursion run in libeksb on the GPU; all
tracks of a call (e.g. both paws) go through ONE batched launch with per-track parameters.
"""
from __future__ import annotations

import numpy as np
import pandas as pd
import torch

from . import _lib, device as dev
from .ensemble_kalman import (ensemble, filtering_pass, kalman_dot, smooth_backward, _cuda,  # noqa: F401 (re-exported like the reference :7)
                              _raise_if_singular, _nonpositive_variance)
from .newton_eks import kalman_newton_recursive  # noqa: F401
from .utils import make_dlc_pandas_index

IMG_WIDTH = 128   # multiview_pca_smoother.py:72


# ---------------------------------------------------------------------------- small helpers
def remove_camera_means(ensemble_stacks, camera_means):
    """(K,T,p) stacks minus a per-column mean (multiview_pca_smoother.py:13-18)."""
    return np.asarray(ensemble_stacks) - np.asarray(camera_means)[None, None, :]


def add_camera_means(ensemble_stacks, camera_means):
    """Inverse of remove_camera_means (:21-26)."""
    return np.asarray(ensemble_stacks) + np.asarray(camera_means)[None, None, :]


class _PCA3:
    """The part of sklearn's PCA the reference uses (:29-31): mean_, components_, transform,
    explained_variance_ratio_.  Component signs are not pinned (outputs C ms are sign-invariant)."""

    def __init__(self, rows, n_comps):
        rows = np.asarray(rows, dtype=np.float64)
        self.mean_ = rows.mean(axis=0)
        _, sv, vt = np.linalg.svd(rows - self.mean_, full_matrices=False)
        self.components_ = vt[:n_comps]
        var = sv ** 2 / max(1, rows.shape[0] - 1)
        self.explained_variance_ratio_ = var[:n_comps] / var.sum()

    def transform(self, rows):
        return (np.asarray(rows) - self.mean_) @ self.components_.T


def pca(S, n_comps):
    p = _PCA3(S, n_comps)
    return p, p.explained_variance_ratio_


def _ensemble_median(stack):
    """stack (M,T,K) host -> (preds (T,K), vars (T,K)) through the CUDA ensemble kernel."""
    d = _cuda()
    avg, var = dev.ensemble(torch.as_tensor(np.ascontiguousarray(stack, dtype=np.float64)).to(d), "median")
    return avg.cpu().numpy(), var.cpu().numpy()


def _good_frames(evars, quantile):
    mx = np.max(evars, axis=1)
    return np.where(mx <= np.percentile(mx, quantile))[0]          # :685-688


def _kf_params(pca_, good_pcs, smooth_param):
    """m0, S0, A, Q, C as set at :721-731 (np.var ddof=0; np.cov of consecutive good rows, ddof=1)."""
    S0 = np.diag(np.var(good_pcs, axis=0))
    Q = smooth_param * np.cov((good_pcs[1:] - good_pcs[:-1]).T)
    return np.zeros(3), S0, np.eye(3), Q, pca_.components_.T


def _run_tracks(tracks, newton=False):
    """tracks: list of dict(y, ev, m0, S0, A, Q, C) of equal T and p.  One batched launch.
    Returns list of (latent (T,n), ym (T,p)); latent = RTS mean, or the Newton-recursive q."""
    d = _cuda()
    B = len(tracks)
    T, p = tracks[0]["y"].shape
    n = tracks[0]["C"].shape[1]
    stk = lambda k: np.stack([t[k] for t in tracks])
    to = lambda a: torch.as_tensor(np.ascontiguousarray(a, dtype=np.float64)).to(d)
    prm = dev.TrackParams.from_numpy(stk("m0"), stk("S0"), stk("C"), stk("A"), stk("Q"), dtype=torch.float64, device=d)
    obs = dev.pack_obs(to(stk("y")), to(stk("ev")))
    sg = dev.sigma(n)
    first = _lib.FIRST_PRIOR_ONLY if newton else _lib.FIRST_PRIOR_UPDATE
    fast = prm.fast_path
    if fast:
        filt, _, status = dev.kf_forward(obs, prm, B, T, first=first)
        if _nonpositive_variance(status):
            fast = False            # a zero ensemble variance: the literal LU kernels have the reference's semantics
    if not fast:
        # (n,p) without a register-resident instantiation (e.g. five cameras, p = 10): runtime-dimension kernels
        filt, status = dev.kf_forward_generic(obs, prm, B, T, first=first)
    if newton:
        lat = dev.unpack_rows(filt, sg, 0, B, T, n).cpu().numpy()
        ym = dev.unpack_rows(dev.project(filt, prm, B, T), p, 0, B, T, p).cpu().numpy()   # y_m = B q (:555, :905)
    else:
        out = (dev.rts_backward(filt, prm, B, T, inplace=True, want_ym=True) if prm.fast_path
               else dev.rts_backward_generic(filt, prm, B, T, want_ym=True))
        status = status | out["status"]
        lat = dev.unpack_rows(out["smth"], sg, 0, B, T, n).cpu().numpy()
        ym = dev.unpack_rows(out["ym"], p, 0, B, T, p).cpu().numpy()  # C ms fused into the RTS kernel (:745)
    _raise_if_singular(status)
    return [(lat[b], ym[b]) for b in range(B)]


# ---------------------------------------------------------------------------- multi-camera
def _stack_cameras(markers_list_cameras):
    """list[camera][model] of DataFrames -> (M,T,2*n_cam); columns 0,1 of every frame (:645-659)."""
    n_cam, M = len(markers_list_cameras), len(markers_list_cameras[0])
    return np.stack([np.concatenate([markers_list_cameras[c][m].to_numpy()[:, :2] for c in range(n_cam)], axis=1)
                     for m in range(M)]).astype(np.float64)


def smooth_session_multi_cam(stack, smooth_param, quantile_keep_pca, newton=False, dtype=torch.float64):
    """All keypoints of a session in one pass, nothing but the result leaving the GPU.

    stack: (M, T, B, p) host array or CUDA tensor -- M ensemble members, T frames, B keypoints (tracks),
    p = 2 * n_cameras columns ordered [cam0_x, cam0_y, cam1_x, ...] (the restack of :645-659).
    Pipeline, every stage a libeksb kernel: ensemble median + variance fused with the restack
    (eksb_ensemble_tiled, :675-682) -> per-track good-frame selection, camera means, PCA(3), S0, Q, C
    (eksb_pca_setup, :684-731; the centring of :702-705 is fused into the forward filter as a per-track
    observation offset) -> filter + RTS + back-projection (or the Newton-recursive filter).
    Returns dict(y=(B,T,p) smoothed observations with the camera means added back (:745,:761-762),
    latent=(B,T,3), means=(B,p)) as host float64 arrays."""
    d = _cuda()
    t = stack if isinstance(stack, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(stack))
    M, T, B, p = t.shape
    if M > 8 or p not in (4, 6, 8):
        return _smooth_session_general(t, smooth_param, quantile_keep_pca, newton)
    preds = t.to(device=d, dtype=dtype).reshape(M, T, B * p).contiguous()
    ws = dev.setup_workspace(B, T, p, dtype, d)                       # the ensemble kernel leaves max_vars in it
    obs = dev.ensemble_tiled(preds, B, p, "median", max_var=dev.setup_max_var_view(ws, B, T, p, dtype))
    prm, means, status = dev.pca_setup(obs, B, T, p, quantile_keep_pca, smooth_param, n=3, center=False,
                                       workspace=ws, mx_ready=True)
    sg = dev.sigma(3)
    first = _lib.FIRST_PRIOR_ONLY if newton else _lib.FIRST_PRIOR_UPDATE
    rws = dev.route_workspace(B, T, 3, p, dtype, d)                   # few keypoints x long session: parallel in time
    filt, _, _ = dev.kf_forward(obs, prm, B, T, first=first, status=status, y_offset=means, workspace=rws)
    if _nonpositive_variance(status):
        # some frame had every ensemble member agree to the last bit: redo the session with the literal LU kernels
        return _smooth_session_general(t, smooth_param, quantile_keep_pca, newton)
    if newton:
        lat = dev.unpack_rows(filt, sg, 0, B, T, 3)
        ym = dev.unpack_rows(dev.project(filt, prm, B, T), p, 0, B, T, p)
    else:
        out = dev.rts_backward(filt, prm, B, T, inplace=True, want_ym=True, status=status, workspace=rws)
        lat = dev.unpack_rows(out["smth"], sg, 0, B, T, 3)
        ym = dev.unpack_rows(out["ym"], p, 0, B, T, p)
    _raise_if_singular(status)                                        # includes "< 3 good frames" from the setup
    y = (ym + means[:, None, :]).cpu().numpy().astype(np.float64)
    return dict(y=y, latent=lat.cpu().numpy().astype(np.float64), means=means.cpu().numpy().astype(np.float64))


def _smooth_session_general(t, smooth_param, quantile_keep_pca, newton):
    """Any ensemble size (<= 64) and any camera count: plain ensemble kernel, the parameter setup of
    multiview_pca_smoother.py:684-731 in host numpy (as the paw path does), runtime-dimension filter / RTS
    kernels where (3, p) has no register-resident instantiation.  Same results, not the fused device pipeline."""
    M, T, B, p = t.shape
    stack = t.detach().cpu().numpy().astype(np.float64) if isinstance(t, torch.Tensor) else np.asarray(t, dtype=np.float64)
    preds, evars = _ensemble_median(stack.reshape(M, T, B * p))
    tracks, means = [], np.empty((B, p))
    for b in range(B):
        yb, evb = preds[:, b * p:(b + 1) * p], evars[:, b * p:(b + 1) * p]
        good = _good_frames(evb, quantile_keep_pca)
        means[b] = yb[good].mean(axis=0)                                   # :698-700
        yc = yb - means[b]
        pca_, _ = pca(yc[good], 3)                                         # :707-708
        m0, S0, A, Q, C = _kf_params(pca_, pca_.transform(yc)[good], smooth_param)
        tracks.append(dict(y=yc, ev=evb, m0=m0, S0=S0, A=A, Q=Q, C=C))
    res = _run_tracks(tracks, newton=newton)
    y = np.stack([r[1] for r in res]) + means[:, None, :]
    return dict(y=y, latent=np.stack([r[0] for r in res]), means=means)


def session_dfs(y, keypoint_names, camera_names):
    """{camera_df: DataFrame (T, 3*B)} with the DLC column index over ALL keypoints (likelihood = NaN),
    the layout scripts/multicam_example.py:118-160 assembles keypoint by keypoint."""
    B, T, _ = y.shape
    out = {}
    for c, name in enumerate(camera_names):
        arr = np.full((T, 3 * B), np.nan)
        arr[:, 0::3] = y[:, :, 2 * c].T
        arr[:, 1::3] = y[:, :, 2 * c + 1].T
        out[name + '_df'] = pd.DataFrame(arr, columns=make_dlc_pandas_index(list(keypoint_names)))
    return out


def ensemble_kalman_smoother_multi_cam(
        markers_list_cameras, keypoint_ensemble, smooth_param, quantile_keep_pca, camera_names):
    """multiview_pca_smoother.py:611-767.  Returns {f'{camera}_df': DataFrame (T,3) [x, y, likelihood=NaN]}."""
    stack = _stack_cameras(markers_list_cameras)[:, :, None, :]
    res = smooth_session_multi_cam(stack, smooth_param, quantile_keep_pca)
    return session_dfs(res["y"], [keypoint_ensemble], camera_names)


def eks_opti_smoother_multi_cam(
        markers_list_cameras, keypoint_ensemble, smooth_param, quantile_keep_pca, camera_names, plot=False):
    """multiview_pca_smoother.py:777-933 with the Newton-recursive filter (:902).  `plot` is accepted
    for signature compatibility; the reference's plot branch dereferences undefined names (:907-913)."""
    stack = _stack_cameras(markers_list_cameras)[:, :, None, :]
    res = smooth_session_multi_cam(stack, smooth_param, quantile_keep_pca, newton=True)
    return session_dfs(res["y"], [keypoint_ensemble], camera_names)


# ---------------------------------------------------------------------------- paw, two async views
def _align_views(left_stack, right_stack, ts_left, ts_right):
    """Right view linearly interpolated onto left timestamps, x flipped (:70-101), all ensemble members
    in one eksb_interp_rows launch.  Left frames before the right camera's first timestamp are skipped and
    the loop stops at the first one beyond its last (:83-86).
    left_stack (M,T_l,6), right_stack (M,T_r,6) -> (left (M,T',4), right (M,T',4)), columns [0,1,3,4]."""
    d = _cuda()
    ts_left, ts_right = np.asarray(ts_left, dtype=np.float64), np.asarray(ts_right, dtype=np.float64)
    beyond = np.nonzero(ts_left > ts_right[-1])[0]
    stop = beyond[0] if beyond.size else ts_left.shape[0]
    keep = np.nonzero(ts_left[:stop] >= ts_right[0])[0]
    cols = [0, 1, 3, 4]
    to = lambda a: torch.as_tensor(np.ascontiguousarray(a, dtype=np.float64)).to(d)
    flip = to([IMG_WIDTH, 0.0, IMG_WIDTH, 0.0])                        # x -> img_width - x (:90-91)
    right = dev.interp_rows(to(ts_right), to(right_stack[:, :, cols]), to(ts_left[keep]), flip)
    return left_stack[:, keep][:, :, cols].astype(np.float64), right.cpu().numpy()


def _paw_setup(markers_list_left_cam, markers_list_right_cam, ts_left, ts_right, smooth_param, quantile_keep_pca):
    left, right = _align_views(np.stack([l.to_numpy() for l in markers_list_left_cam]),
                               np.stack([r.to_numpy() for r in markers_list_right_cam]), ts_left, ts_right)
    # one ensemble launch over both views: columns [left l_x l_y r_x r_y | right l_x l_y r_x r_y]  (:115-118)
    stack = np.concatenate([left, right], axis=2)
    preds, evars = _ensemble_median(stack)
    good = _good_frames(evars, quantile_keep_pca)                      # max over both views (:123-126)
    lp, rp, lv, rv = preds[:, :4], preds[:, 4:], evars[:, :4], evars[:, 4:]
    y_paw = [np.hstack([lp[:, :2], rp[:, :2]]), np.hstack([lp[:, 2:4], rp[:, 2:4]])]     # :154-170
    ev_paw = [np.hstack([lv[:, :2], rv[:, :2]]), np.hstack([lv[:, 2:4], rv[:, 2:4]])]
    stacked = np.empty((2 * good.shape[0], 4))                         # paws interleaved on good frames (:135-151)
    stacked[0::2] = y_paw[0][good]
    stacked[1::2] = y_paw[1][good]
    means = stacked.mean(axis=0)                                       # :173-177
    pca_, _ = pca(stacked - means, 3)                                  # shared by both paws (:189-191)
    tracks = []
    for k in range(2):
        y = y_paw[k] - means
        m0, S0, A, Q, C = _kf_params(pca_, pca_.transform(y)[good], smooth_param)
        tracks.append(dict(y=y, ev=ev_paw[k], m0=m0, S0=S0, A=A, Q=Q, C=C))
    return tracks, means


def _paw_dfs(ym_left_paw, ym_right_paw, means, keypoint_names):
    a, b = ym_left_paw + means, ym_right_paw + means                   # add_camera_means (:282)
    nan = np.full(a.shape[0], np.nan)
    # left view: paw_l, paw_r.  right view: paws swapped back and x un-flipped (:296-320)
    left = np.column_stack([a[:, 0], a[:, 1], nan, b[:, 0], b[:, 1], nan])
    right = np.column_stack([IMG_WIDTH - b[:, 2], b[:, 3], nan, IMG_WIDTH - a[:, 2], a[:, 3], nan])
    idx = make_dlc_pandas_index(keypoint_names)
    return {'left_df': pd.DataFrame(left, columns=idx), 'right_df': pd.DataFrame(right, columns=idx)}


def ensemble_kalman_smoother_paw_asynchronous(
        markers_list_left_cam, markers_list_right_cam, timestamps_left_cam,
        timestamps_right_cam, keypoint_names, smooth_param, quantile_keep_pca):
    """multiview_pca_smoother.py:34-322.  Returns {'left_df', 'right_df'} (T',6)."""
    tracks, means = _paw_setup(markers_list_left_cam, markers_list_right_cam, timestamps_left_cam,
                               timestamps_right_cam, smooth_param, quantile_keep_pca)
    (_, ym_l), (_, ym_r) = _run_tracks(tracks)
    return _paw_dfs(ym_l, ym_r, means, keypoint_names)


def eks_opti_smoother_paw_asynchronous(
        markers_list_left_cam, markers_list_right_cam, timestamps_left_cam,
        timestamps_right_cam, keypoint_names, smooth_param, quantile_keep_pca):
    """multiview_pca_smoother.py:326-608: same setup, Newton-recursive filter (:552)."""
    tracks, means = _paw_setup(markers_list_left_cam, markers_list_right_cam, timestamps_left_cam,
                               timestamps_right_cam, smooth_param, quantile_keep_pca)
    (_, ym_l), (_, ym_r) = _run_tracks(tracks, newton=True)
    return _paw_dfs(ym_l, ym_r, means, keypoint_names)
