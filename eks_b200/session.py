"""Whole-session driver: the work of the reference's scripts/multicam_example.py:81-160 (load every ensemble
member's CSV, smooth each body part across cameras, write one eks.csv) with the data path native end to end:
CSV bodies parsed by libeksb into one pinned (M, T, cols) array -> one H2D copy -> column gather (the restack of
multiview_pca_smoother.py:645-659) -> ensemble / parameter setup / filter / RTS kernels for ALL body parts in
one batch -> one D2H copy -> CSV written by libeksb.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from . import io
from .ensemble_kalman import _cuda
from .multiview_pca_smoother import smooth_session_multi_cam

TRACKER = "ensemble-kalman_tracker"


def smooth_multicam_csvs(csv_paths, bodypart_list, camera_names, smooth_param, quantile_keep_pca, save_path=None,
                         eks_version="standard"):
    """csv_paths: one DLC/LP prediction CSV per ensemble member, every camera view in the same file with
    keypoints named f'{bodypart}_{camera}' (scripts/multicam_example.py:107-118).
    Returns dict(data (T, cols) float64 -- the eks.csv body: smoothed x/y, likelihood 1.0, untouched columns NaN
    (:95-103), header, index, y (B,T,p), latent (B,T,3)); writes save_path when given."""
    d = _cuda()
    ens = io.load_ensemble(csv_paths, device=d)
    cols = io.multicam_columns(ens["bodyparts"], ens["coords"], bodypart_list, camera_names)   # (B,p)
    B, p = cols.shape
    M, T, _ = ens["stack"].shape
    idx = torch.as_tensor(cols.reshape(-1), device=d)
    stack = ens["stack"].index_select(2, idx).reshape(M, T, B, p)
    res = smooth_session_multi_cam(stack, smooth_param, quantile_keep_pca, newton=(eks_version == "opti"))
    K = len(ens["coords"])
    body = np.full((T, K), np.nan)
    body[:, [i for i, c in enumerate(ens["coords"]) if c == "likelihood"]] = 1.0
    body[:, cols.reshape(-1)] = res["y"].transpose(1, 0, 2).reshape(T, B * p)
    lines = ens["header"].split("\n")
    lines[0] = ",".join([lines[0].split(",")[0]] + [TRACKER] * K)                              # scorer level (:96)
    header = "\n".join(lines)
    if save_path is not None:
        os.makedirs(os.path.dirname(os.path.abspath(save_path)), exist_ok=True)
        io.write_dlc_csv(save_path, header, ens["index"], body)
    return dict(data=body, header=header, index=ens["index"], y=res["y"], latent=res["latent"])
