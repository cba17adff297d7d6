"""Drop-in for ``eks/utils.py``: DLC-style column index helpers (host-side pandas only)."""
import pandas as pd


def make_dlc_pandas_index(keypoint_names):
    """3-level column index (scorer, bodyparts, coords) used by every smoother output (eks/utils.py:4-10)."""
    return pd.MultiIndex.from_product(
        [["ensemble-kalman_tracker"], list(keypoint_names), ["x", "y", "likelihood"]],
        names=["scorer", "bodyparts", "coords"])


def convert_lp_dlc(df_lp, keypoint_names, model_name=None):
    """Flatten a Lightning-Pose / DLC 3-row-header frame into '{keypoint}_{x|y|likelihood}' columns
    (eks/utils.py:13-22)."""
    cols = {}
    for kp in keypoint_names:
        for coord in ("x", "y", "likelihood"):
            key = (kp, coord) if model_name is None else (model_name, kp, coord)
            cols[f"{kp}_{coord}"] = df_lp.loc[:, key]
    return pd.DataFrame(cols, index=df_lp.index)
