"""Drop-in for ``eks/pupil_smoother.py`` (erialc-cal/eks): pupil diameter / centre-of-mass
smoother and its Newton ("opti") twin.  Geometry helpers and DataFrame assembly are host
numpy/pandas (O(T) bookkeeping that only defines m0, S0, Q and the means); the ensemble
reduction, Kalman filter, RTS smoother and Newton recursion run in libeksb on the GPU.
"""
from __future__ import annotations

import warnings

import numpy as np
import pandas as pd

from .ensemble_kalman import ensemble, filtering_pass, kalman_dot, smooth_backward  # noqa: F401 (re-exported, reference :7)
from .multiview_pca_smoother import _run_tracks
from .newton_eks import kalman_newton_recursive  # noqa: F401
from .utils import make_dlc_pandas_index

KEYS = ['pupil_top_r_x', 'pupil_top_r_y', 'pupil_bottom_r_x', 'pupil_bottom_r_y',
        'pupil_right_r_x', 'pupil_right_r_y', 'pupil_left_r_x', 'pupil_left_r_y']      # pupil_smoother.py:102-103


def _xy(dlc, point):
    return np.column_stack([np.asarray(dlc[f'pupil_{point}_r_x'], dtype=np.float64),
                            np.asarray(dlc[f'pupil_{point}_r_y'], dtype=np.float64)])


def get_pupil_location(dlc):
    """Pupil centre (T,2): median-of-medians of the four edge points (pupil_smoother.py:14-41):
    x from {top,bottom} (NaN-tolerant) and {right,left} (both needed); y the other way round."""
    t, b, l, r = (_xy(dlc, k) for k in ('top', 'bottom', 'left', 'right'))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", category=RuntimeWarning)
        x1 = np.nanmedian(np.column_stack([t[:, 0], b[:, 0]]), axis=1)
        x2 = np.median(np.column_stack([r[:, 0], l[:, 0]]), axis=1)
        y1 = np.median(np.column_stack([t[:, 1], b[:, 1]]), axis=1)
        y2 = np.nanmedian(np.column_stack([r[:, 1], l[:, 1]]), axis=1)
        return np.column_stack([np.nanmedian(np.column_stack([x1, x2]), axis=1),
                                np.nanmedian(np.column_stack([y1, y2]), axis=1)])


def get_pupil_diameter(dlc):
    """nanmedian of six diameter estimates: top-bottom, left-right and the four adjacent-edge
    chords scaled by sqrt(2) (circle assumption)  (pupil_smoother.py:44-68)."""
    t, b, l, r = (_xy(dlc, k) for k in ('top', 'bottom', 'left', 'right'))
    est = [np.linalg.norm(t - b, axis=1), np.linalg.norm(l - r, axis=1)]
    est += [np.linalg.norm(u - v, axis=1) * 2 ** 0.5 for u, v in ((t, l), (t, r), (b, l), (b, r))]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", category=RuntimeWarning)
        return np.nanmedian(np.asarray(est), axis=0)


def add_mean_to_array(pred_arr, keys, mean_x, mean_y):
    """{key: column + mean_x if 'x' in key else + mean_y}  (pupil_smoother.py:71-79)."""
    return {key: pred_arr[:, i] + (mean_x if 'x' in key else mean_y) for i, key in enumerate(keys)}


C_PUPIL = np.asarray([[0, 1, 0], [-.5, 0, 1], [0, 1, 0], [.5, 0, 1],
                      [.5, 1, 0], [0, 0, 1], [-.5, 1, 0], [0, 0, 1]], dtype=np.float64)  # :150-153


def _setup(markers_list, A):
    preds, evars, _, mean_dict, _, _ = ensemble(markers_list, KEYS)       # CUDA ensemble kernel (:104)
    loc = get_pupil_location(mean_dict)
    diam = get_pupil_diameter(mean_dict)
    mean_x, mean_y = np.mean(loc[:, 0]), np.mean(loc[:, 1])
    var = np.asarray([np.var(diam), np.var(loc[:, 0] - mean_x), np.var(loc[:, 1] - mean_y)])
    A = np.asarray(A, dtype=np.float64)
    y = preds.copy()
    y[:, 0::2] -= mean_x                                                   # :158-163
    y[:, 1::2] -= mean_y
    trk = dict(y=y, ev=evars, m0=np.asarray([np.mean(diam), 0.0, 0.0]), S0=np.diag(var), A=A,
               Q=np.diag(var * (1.0 - np.diag(A) ** 2)), C=C_PUPIL)        # :130-153
    return trk, mean_x, mean_y


def _frames(ym, latents, mean_x, mean_y, keypoint_names, tracker_name):
    cols = add_mean_to_array(ym, KEYS, mean_x, mean_y)
    nan = np.full(ym.shape[0], np.nan)
    arr = []
    for point in ('top', 'right', 'bottom', 'left'):                      # :199-202
        arr += [cols[f'pupil_{point}_r_x'], cols[f'pupil_{point}_r_y'], nan]
    markers_df = pd.DataFrame(np.column_stack(arr), columns=make_dlc_pandas_index(keypoint_names))
    lat = np.column_stack([latents[:, 0], latents[:, 1] + mean_x, latents[:, 2] + mean_y])
    idx = pd.MultiIndex.from_arrays([[tracker_name] * 3, ['diameter', 'com_x', 'com_y']], names=('scorer', 'latent'))
    return {'markers_df': markers_df, 'latents_df': pd.DataFrame(lat, columns=idx)}


def ensemble_kalman_smoother_pupil(markers_list, keypoint_names, tracker_name, state_transition_matrix):
    """pupil_smoother.py:82-223.  Returns {'markers_df' (T,12), 'latents_df' (T,3)}."""
    trk, mean_x, mean_y = _setup(markers_list, state_transition_matrix)
    (ms, ym), = _run_tracks([trk])
    return _frames(ym, ms, mean_x, mean_y, keypoint_names, tracker_name)


def eks_opti_smoother_pupil(markers_list, keypoint_names, tracker_name, state_transition_matrix, plot=True):
    """pupil_smoother.py:227-364.  Like the reference it ignores `state_transition_matrix` and uses
    A = 0.99 I (:301-305).  latents_df are the Newton-recursive latents (pinned by the shipped
    opti_eks_latents.csv).  markers_df with the reference's default plot=True: the reference then overwrites q with
    the latents CSV of the STANDARD smoother (:325-329; its shipped kalman_smoothed_latents.csv is the standard
    smoother at A = 0.99 I with the means already added to com_x / com_y) and returns B q + means (:332-346) --
    reproduced here by running the standard smoother instead of reading a hard-coded /Users path, and pinned by
    the shipped opti_eks_pupil_traces.csv.  With plot=False the reference raises UnboundLocalError (q is never
    assigned); here that case returns markers_df = B q_newton + means.  Nothing is plotted either way."""
    A = np.diag([0.99, 0.99, 0.99])
    trk, mean_x, mean_y = _setup(markers_list, A)
    if plot:
        (q, _), = _run_tracks([trk], newton=True)
        (ms, _), = _run_tracks([trk])
        std_latents = np.column_stack([ms[:, 0], ms[:, 1] + mean_x, ms[:, 2] + mean_y])     # what the CSV holds
        return _frames(std_latents @ C_PUPIL.T, q, mean_x, mean_y, keypoint_names, tracker_name)
    (q, ym), = _run_tracks([trk], newton=True)
    return _frames(ym, q, mean_x, mean_y, keypoint_names, tracker_name)
