"""DLC / Lightning-Pose CSV reader + writer against pandas, the reference's I/O
(scripts/multicam_example.py:85,139).  Host-only library functions: no GPU needed."""
import numpy as np
import pandas as pd

from eks_b200 import io


def _dlc_frame(T, keypoints, seed=0, scorer="heatmap_tracker"):
    rng = np.random.default_rng(seed)
    cols = pd.MultiIndex.from_product([[scorer], keypoints, ["x", "y", "likelihood"]], names=["scorer", "bodyparts", "coords"])
    data = rng.uniform(0, 400, size=(T, len(cols)))
    data[:, 2::3] = rng.uniform(0, 1, size=(T, len(keypoints)))
    data[3, 0] = 1e-7; data[4, 1] = 123456789012345678.0; data[5, 3] = -0.0001; data[6, 4] = 5.0; data[7, 0] = 1e16; data[8, 1] = 1e-4
    return pd.DataFrame(data, columns=cols)


def test_reader_matches_pandas_round_trip_parser(tmp_path):
    df = _dlc_frame(1500, ["paw1_top", "paw1_bot", "nose_top", "nose_bot"])
    df.iloc[10, 5] = np.nan
    path = tmp_path / "preds.csv"
    df.to_csv(path)
    got = io.read_dlc_csv(str(path))
    exact = pd.read_csv(path, header=[0, 1, 2], index_col=0, float_precision="round_trip")
    assert np.array_equal(got["data"], exact.to_numpy(), equal_nan=True)            # correctly rounded: bit-exact
    assert np.array_equal(got["data"], df.to_numpy(), equal_nan=True)               # ... and round-trips to_csv
    # the reference's own call uses pandas' default converter, which is not correctly rounded (it drops
    # digits beyond ~15 significant): agreement to 1e-12 relative
    default = pd.read_csv(path, header=[0, 1, 2], index_col=0).to_numpy()
    ok = np.isnan(default) | (np.abs(default - got["data"]) <= 1e-12 * np.abs(default))
    assert ok.all()
    assert np.array_equal(got["index"], df.index.to_numpy())
    assert got["bodyparts"] == [c[1] for c in df.columns] and got["coords"] == [c[2] for c in df.columns]
    assert got["scorer"][0] == "heatmap_tracker"


def test_writer_text_is_identical_to_to_csv(tmp_path):
    df = _dlc_frame(300, ["a_top", "a_bot"], seed=3)
    df.iloc[::7, 2] = np.nan                                                          # likelihood = NaN like the smoother outputs
    ref, ours = tmp_path / "ref.csv", tmp_path / "ours.csv"
    df.to_csv(ref)
    r = io.read_dlc_csv(str(ref))
    io.write_dlc_csv(str(ours), r["header"], r["index"], r["data"])
    assert open(ours).read() == open(ref).read()


def test_load_ensemble_stacks_members(tmp_path):
    paths = []
    for m in range(3):
        df = _dlc_frame(64, ["k_top", "k_bot"], seed=m, scorer=f"model{m}")
        p = tmp_path / f"m{m}.csv"
        df.to_csv(p)
        paths.append(str(p))
    ens = io.load_ensemble(paths, device=None, pin=False)
    assert tuple(ens["stack"].shape) == (3, 64, 6) and ens["scorer"] == ["model0", "model1", "model2"]
    cols = io.multicam_columns(ens["bodyparts"], ens["coords"], ["k"], ["top", "bot"])
    assert cols.tolist() == [[0, 1, 3, 4]]
    for m in range(3):
        want = pd.read_csv(paths[m], header=[0, 1, 2], index_col=0, float_precision="round_trip").to_numpy()
        assert np.array_equal(ens["stack"][m].numpy(), want)
