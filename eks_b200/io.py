"""On-disk format either side of the hot path: DeepLabCut / Lightning-Pose prediction CSVs.

The reference loads every ensemble member with ``pd.read_csv(path, header=[0, 1, 2], index_col=0)`` and
restacks DataFrames frame by frame (scripts/multicam_example.py:81-103, multiview_pca_smoother.py:645-659);
results go back out through ``DataFrame.to_csv`` (:139,:157).  Here the numeric body is parsed natively
(libeksb ``eksb_csv_dlc_read_host``, multi-threaded) straight into one pinned ``(M, T, cols)`` array that is
copied to the GPU once, and written with ``eksb_csv_dlc_write_host`` (same text as ``to_csv``).
"""
from __future__ import annotations

import ctypes
import os

import numpy as np
import torch

from ._lib import check, lib


def _header_levels(header_text):
    """3 raw header lines -> (scorer, bodyparts, coords) lists, index column dropped."""
    rows = [ln.rstrip("\r").split(",")[1:] for ln in header_text.split("\n")[:3]]
    return rows[0], rows[1], rows[2]


def read_dlc_csv(path, out=None, n_threads=0):
    """-> dict(data (T, cols) float64 ndarray, index (T,) int64, header (raw 3 lines), scorer, bodyparts, coords).
    `out`: optional preallocated (T, cols) float64 C-contiguous array (e.g. a view of pinned memory)."""
    L = lib()
    bpath = os.fsencode(path)
    n_rows, n_cols = ctypes.c_long(), ctypes.c_int()
    check(L.eksb_csv_dlc_shape_host(bpath, ctypes.addressof(n_rows), ctypes.addressof(n_cols)), "eksb_csv_dlc_shape_host")
    T, K = n_rows.value, n_cols.value
    data = np.empty((T, K), dtype=np.float64) if out is None else out
    if data.shape != (T, K) or data.dtype != np.float64 or not data.flags.c_contiguous:
        raise ValueError(f"out must be a C-contiguous float64 array of shape {(T, K)}")
    index = np.empty(T, dtype=np.int64)
    hdr = ctypes.create_string_buffer(1 << 20)
    check(L.eksb_csv_dlc_read_host(bpath, data.ctypes.data, T, K, index.ctypes.data, ctypes.addressof(hdr), len(hdr),
                                   int(n_threads)), "eksb_csv_dlc_read_host")
    header = hdr.value.decode()
    scorer, bodyparts, coords = _header_levels(header)
    return dict(data=data, index=index, header=header, scorer=scorer, bodyparts=bodyparts, coords=coords)


def write_dlc_csv(path, header, index, data):
    """Write (T, cols) float64 `data` under the 3 raw header lines; text identical to DataFrame.to_csv."""
    data = np.ascontiguousarray(data, dtype=np.float64)
    index = np.ascontiguousarray(index, dtype=np.int64)
    check(lib().eksb_csv_dlc_write_host(os.fsencode(path), header.encode(), index.ctypes.data, data.ctypes.data,
                                        data.shape[0], data.shape[1]), "eksb_csv_dlc_write_host")


def load_ensemble(csv_paths, device=None, pin=True):
    """M prediction CSVs of one session -> dict(stack (M, T, cols) tensor on `device` (pinned host if None),
    bodyparts, coords, scorer, index, header).  All files must share the column layout."""
    metas, stack = [], None
    for m, path in enumerate(csv_paths):
        if stack is None:
            first = read_dlc_csv(path)
            T, K = first["data"].shape
            stack = torch.empty((len(csv_paths), T, K), dtype=torch.float64, pin_memory=pin and torch.cuda.is_available())
            stack[0].copy_(torch.from_numpy(first["data"]))
            metas.append(first)
        else:
            metas.append(read_dlc_csv(path, out=stack[m].numpy()))
            if metas[-1]["bodyparts"] != metas[0]["bodyparts"] or metas[-1]["coords"] != metas[0]["coords"]:
                raise ValueError(f"{path}: column layout differs from {csv_paths[0]}")
    meta = metas[0]
    if device is not None:
        stack = stack.to(device, non_blocking=True)
    return dict(stack=stack, bodyparts=meta["bodyparts"], coords=meta["coords"], scorer=[m_["scorer"][0] for m_ in metas],
                index=meta["index"], header=meta["header"])


def multicam_columns(bodyparts, coords, keypoints, camera_names):
    """Column indices (B, p) into a DLC body for keypoints named f'{keypoint}_{camera}' (the mirror-mouse
    convention, scripts/multicam_example.py:110-118): p = [cam0_x, cam0_y, cam1_x, ...]."""
    pos = {(b, c): i for i, (b, c) in enumerate(zip(bodyparts, coords))}
    return np.array([[pos[(f"{k}_{cam}", xy)] for cam in camera_names for xy in ("x", "y")] for k in keypoints])
