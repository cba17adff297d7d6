"""Randomised parity sweep (developer probe): many small (B, T, n, p, dtype, A-mode) cases through the C ABI,
pipelined kernels vs the numpy oracle; extreme ensemble variances (1e-4 .. 1e3), ragged B and T, T in {1,2,3,..}.
Metric and bars are the tests' own (tests/conftest.py rel: per component, scaled by its largest magnitude over time;
1e-3 in fp32 = north_star's bar, 1e-8 in fp64)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from eks_b200 import device as dev
from oracle import restated as o
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import rel

rng = np.random.default_rng(int(os.environ.get("SEED", "0")))
d = torch.device("cuda", 0)
worst = {}
worst_n = {}
nfail = 0
for it in range(int(os.environ.get("ITERS", "150"))):
    n, p = [(3, 4), (3, 6), (3, 8), (2, 4), (4, 6), (6, 8)][rng.integers(6)]
    B = int(rng.choice([1, 2, 31, 32, 33, 64, 70, 150]))
    T = int(rng.choice([1, 2, 3, 4, 5, 6, 7, 8, 9, 11, 12, 13, 17, 24, 25, 48, 49, 97, 200, 333]))
    dt, tol = (torch.float64, 1e-8) if rng.random() < 0.5 else (torch.float32, 1e-3)
    a_diag = rng.random() < 0.7
    C = np.stack([np.linalg.qr(rng.normal(size=(p, n)))[0] for _ in range(B)])
    A = np.diag(rng.uniform(0.9, 1.0, n)) if a_diag else np.eye(n) * 0.95 + 0.02 * rng.normal(size=(n, n))
    Q = np.diag(10.0 ** rng.uniform(-4, -1, n)); Q = Q + 1e-5
    S0 = np.diag(rng.uniform(0.1, 5, n)); m0 = rng.normal(size=n)
    x = np.cumsum(rng.normal(size=(B, T, n)) * 0.2, axis=1)
    lo, hi = (-4, 3) if dt == torch.float64 else (-3, 2)
    ev = 10.0 ** rng.uniform(lo, hi, size=(B, T, p)) if rng.random() < 0.5 else np.exp(rng.normal(np.log(0.05), 1.0, size=(B, T, p)))
    y = np.einsum("bpn,btn->btp", C, x) + np.sqrt(ev) * rng.normal(size=(B, T, p))
    prm = dev.TrackParams.from_numpy(m0, S0, C, A, Q, dtype=dt, device=d)
    obs = dev.pack_obs(torch.as_tensor(y, dtype=dt).to(d), torch.as_tensor(ev, dtype=dt).to(d))
    filt, _, st = dev.kf_forward(obs, prm, B, T)
    mfV = dev.unpack_rows(filt, dev.sigma(n), 0, B, T, dev.sigma(n)).cpu().numpy().astype(np.float64)
    out = dev.rts_backward(filt, prm, B, T, inplace=False, want_ym=True, want_yv=bool(rng.random() < 0.5))
    sm = dev.unpack_rows(out["smth"], dev.sigma(n), 0, B, T, dev.sigma(n)).cpu().numpy().astype(np.float64)
    ym = dev.unpack_rows(out["ym"], p, 0, B, T, p).cpu().numpy().astype(np.float64)
    assert int(st.sum()) == 0 and int(out["status"].sum()) == 0
    iu = np.triu_indices(n)
    yq, evq = (y, ev) if dt == torch.float64 else (y.astype(np.float32).astype(np.float64), ev.astype(np.float32).astype(np.float64))
    for b in {0, B - 1, B // 2}:
        mf, Vf, S = o.filtering_pass(yq[b], m0, S0, C[b], np.eye(p), A, Q, evq[b])
        ms, Vs, _ = o.smooth_backward(yq[b], mf, Vf, S, A, Q, C[b])
        errs = dict(mf=rel(mfV[b][:, :n], mf), Vf=rel(mfV[b][:, n:], Vf[:, iu[0], iu[1]]), ms=rel(sm[b][:, :n], ms),
                    Vs=rel(sm[b][:, n:], Vs[:, iu[0], iu[1]]), ym=rel(ym[b], ms @ C[b].T))
        for k, e in errs.items():
            key = (str(dt)[6:], k)
            worst[key] = max(worst.get(key, 0.0), e)
            kn = (str(dt)[6:], f"n={n}", "Vs/Vf" if k in ("Vs", "Vf") else "means")
            worst_n[kn] = max(worst_n.get(kn, 0.0), e)
            if not (e < tol):
                nfail += 1
                print("FAIL", dict(n=n, p=p, B=B, T=T, dt=str(dt), a_diag=a_diag, b=b), k, e, flush=True)
print("worst relative errors:", {k: float(f"{v:.3g}") for k, v in sorted(worst.items())})
print("worst by state dimension:", {" ".join(k): float(f"{v:.3g}") for k, v in sorted(worst_n.items())})
print("failures (fp32 bar 1e-3, fp64 1e-8, per component):", nfail)
