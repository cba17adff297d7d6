"""CPU: the C-ABI library loads and exports every symbol include/eksb.h declares (no compute)."""
import ctypes
import os
import subprocess

import pytest

from eks_b200 import _lib


def test_header_parses_and_every_symbol_is_exported():
    protos = _lib.parse_header()
    assert len(protos) >= 20
    assert os.path.exists(_lib.LIB_PATH), "build libeksb.so first (make -C eks_b200/csrc -j8)"
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in protos:
        assert hasattr(raw, name), f"{name} declared in include/eksb.h but not exported"
    out = subprocess.check_output(["nm", "-D", "--defined-only", _lib.LIB_PATH], text=True)
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert {n for n in exported if n.startswith("eksb_")} == set(protos), "exports and header differ"


def test_non_compute_entry_points():
    lib = _lib.lib()
    assert lib.eksb_version() >= 100
    assert lib.eksb_num_tiles(1) == 1 and lib.eksb_num_tiles(32) == 1 and lib.eksb_num_tiles(33) == 2
    for n, p in _lib.SUPPORTED_DIMS:
        assert lib.eksb_supported_dims(n, p) == 1
    # (n,p) outside that list, e.g. (5,4) or five cameras' (3,10), are served by the runtime-dimension kernels
    # eksb_kf_forward_generic / eksb_rts_backward_generic (tests/test_gpu_round2.py::test_shapes_without_a_template_instantiation)
    assert lib.eksb_route_workspace_bytes(2048, 100000, 3, 4, 8) > 0 and lib.eksb_route_workspace_bytes(16384, 100000, 3, 4, 8) == 0
    # time-split NLL sweep: few tiles x long tracks are routed, a batch that already fills the GPU with 8-chain CTAs is not,
    # short tracks are not, shapes without the pipelined NLL kernel are not
    assert lib.eksb_nll_workspace_bytes(32, 100000, 3, 4, 16, 8) > 0 and lib.eksb_nll_workspace_bytes(8192, 20000, 3, 4, 0, 8) > 0
    assert lib.eksb_nll_workspace_bytes(8192, 20000, 3, 4, 16, 8) == 0 and lib.eksb_nll_workspace_bytes(32, 4000, 3, 4, 16, 8) == 0
    assert lib.eksb_nll_workspace_bytes(32, 100000, 4, 6, 16, 8) == 0 and lib.eksb_nll_workspace_bytes(32, 100000, 3, 4, 16, 3) == 0
    assert lib.eksb_kf_nll_ws_f64(None, None, None, None, None, None, None, 31, 4, 100, 3, 4, 0, 0, None, 0, None, None, None, 0, None) == -2
    # argument validation happens before any CUDA call
    assert lib.eksb_ensemble_f64(None, 5, 10, 4, 0, None, None, None) == -2
    assert b"bad argument" in lib.eksb_last_error_string()


def test_product_never_imports_the_oracle():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for dirpath, _, files in os.walk(os.path.join(root, "eks_b200")):
        for f in files:
            if f.endswith(".py"):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text and "hostsim" not in text, f
