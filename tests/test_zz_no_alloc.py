"""GPU: eksb_kf_rts_scan_static_f64, the scan with its workspace in a static device array inside the library (no
allocation in the call, no counterpart in the reference).  Called
through the C ABI and checked against eksb_kf_rts_scan_f64 with a caller-allocated workspace, on two calls in a row
with different inputs (the array keeps the previous call's contents).  Collected last on purpose
(file name)."""
import ctypes

import numpy as np
import pytest

from conftest import rel

pytestmark = pytest.mark.gpu

STATIC_BYTES = 32 << 20        # EKSB_SCAN_STATIC_SCRATCH_BYTES (include/eksb.h)


def _inputs(B, T, seed, d):
    import torch
    from eks_b200 import device as dev
    rng = np.random.default_rng(seed)
    p, n = 4, 3
    C = np.linalg.qr(rng.normal(size=(B, p, n)))[0]
    x = np.cumsum(rng.normal(size=(B, T, n)) * 0.1, axis=1)
    ev = np.exp(rng.normal(np.log(0.1), 0.8, size=(B, T, p)))
    y = np.einsum("bpn,btn->btp", C, x) + np.sqrt(5 * ev) * rng.normal(size=(B, T, p))
    prm = dev.TrackParams.from_numpy(np.zeros(n), np.eye(n), C, np.diag([1.0, 0.999, 0.99]), 1e-2 * np.eye(n), device=d)
    obs = dev.pack_obs(torch.as_tensor(y).to(d), torch.as_tensor(ev).to(d))
    return prm, obs


def _call_static(lib, obs, prm, B, T, chunk, filt, smth, ym, status, stream):
    return lib.eksb_kf_rts_scan_static_f64(obs.data_ptr(), prm.m0.data_ptr(), prm.S0.data_ptr(), prm.C.data_ptr(),
                                           prm.A.data_ptr(), prm.Q.data_ptr(), prm.shared_mask, B, T, prm.n, prm.p,
                                           prm.a_mode, chunk, filt.data_ptr(), smth.data_ptr(), ym.data_ptr(),
                                           status.data_ptr(), stream)


def test_static_scratch_scan_equals_the_workspace_scan_twice_in_a_row():
    import torch
    from eks_b200 import device as dev, _lib
    lib = _lib.lib()
    d = torch.device("cuda", 0)
    torch.cuda.set_device(d)
    stream = torch.cuda.current_stream(d).cuda_stream
    # two calls in a row, different shapes and chunkings: the second one finds the first one's summaries, cuts and
    # flags in the array
    for B, T, chunk, seed in ((70, 6000, 0, 1), (40, 9000, 500, 2)):
        prm, obs = _inputs(B, T, seed, d)
        want = dev.kf_rts_scan(obs, prm, B, T, chunk=chunk, inplace=False)          # caller-allocated workspace
        assert int(want["status"].sum()) == 0
        assert lib.eksb_scan_workspace_bytes(B, T, 3, chunk, 8) <= STATIC_BYTES
        filt = dev.alloc_records(B, T, 9, torch.float64, d)
        smth = torch.empty_like(filt)
        ym = dev.alloc_records(B, T, 4, torch.float64, d)
        status = torch.zeros(B, dtype=torch.int32, device=d)
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info(d)
        rc = _call_static(lib, obs, prm, B, T, chunk, filt, smth, ym, status, stream)
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info(d)
        assert rc == 0, lib.eksb_last_error_string().decode()
        assert int(status.sum()) == 0
        if seed == 2:                      # set-up (module load, symbol lookup) happened in the first call at the latest
            assert free_after >= free_before, "the no-allocation entry point took device memory"
        for got, ref, rows in ((filt, want["filt"], 9), (smth, want["smth"], 9), (ym, want["ym"], 4)):
            a = dev.unpack_rows(got, rows, 0, B, T, rows).cpu().numpy()
            b = dev.unpack_rows(ref, rows, 0, B, T, rows).cpu().numpy()
            assert rel(a, b, time_axis=1) < 1e-9      # the f64 tolerance of the other scan tests (same kernels: ~0)


def test_static_scratch_scan_rejects_requests_beyond_its_array():
    import torch
    from eks_b200 import device as dev, _lib
    lib = _lib.lib()
    d = torch.device("cuda", 0)
    B, T, chunk = 4096, 100_000, 64                    # 128 tiles x 1563 chunks x 16 KiB: far beyond 32 MiB
    assert lib.eksb_scan_workspace_bytes(B, T, 3, chunk, 8) > STATIC_BYTES
    one = torch.zeros(64, dtype=torch.float64, device=d)     # never dereferenced: the size check comes first
    st = torch.zeros(1, dtype=torch.int32, device=d)
    rc = lib.eksb_kf_rts_scan_static_f64(one.data_ptr(), one.data_ptr(), one.data_ptr(), one.data_ptr(), one.data_ptr(),
                                         one.data_ptr(), 31, B, T, 3, 4, 0, chunk, one.data_ptr(), one.data_ptr(),
                                         one.data_ptr(), st.data_ptr(), None)
    assert rc == _lib_earg()
    msg = lib.eksb_last_error_string().decode()
    assert "static scratch" in msg and str(STATIC_BYTES) in msg


def _lib_earg():
    return -2                                          # EKSB_EARG
