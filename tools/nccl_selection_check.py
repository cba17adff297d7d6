"""Multi-GPU check of the path's one collective: smoothing-parameter grid search with tracks sharded
over the ranks, per-GPU NLL sums from libeksb, all-reduce(SUM) over NCCL, common argmin.  Every rank also
evaluates the whole batch alone and compares.  Launch: torchrun --nproc-per-node N tools/nccl_selection_check.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from eks_b200 import device as dev, selection as sel

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
d = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=d)
B, T, p, n = 4099, 3000, 4, 3
rng = np.random.default_rng(0)                                  # same problem on every rank
C = np.linalg.qr(rng.normal(size=(B, p, n)))[0]
x = np.cumsum(rng.normal(size=(B, T, n)) * 0.1, axis=1)
ev = np.exp(rng.normal(np.log(0.05), 0.8, size=(B, T, p)))
y = np.einsum("bpn,btn->btp", C, x) + np.sqrt(ev) * rng.normal(size=(B, T, p))
grid = [1e-3, 1e-2, 1e-1, 1.0, 10.0]
Q0 = 0.01 * np.eye(n)
def run(lo, hi):
    prm = dev.TrackParams.from_numpy(np.zeros(n), np.eye(n), C[lo:hi], np.eye(n), Q0, dtype=torch.float64, device=d)
    obs = dev.pack_obs(torch.as_tensor(y[lo:hi]).to(d), torch.as_tensor(ev[lo:hi]).to(d))
    return obs, prm
lo, hi = sel.shard_bounds(B, rank, world)
obs, prm = run(lo, hi)
best, tot = sel.select_smooth_param(obs, prm, hi - lo, T, grid)          # NCCL all-reduce inside
obs_all, prm_all = run(0, B)
nll, _ = sel.nll_grid(obs_all, prm_all, B, T, grid)
ref = sel.nll_sums(nll).cpu().numpy()
err = np.abs(tot - ref).max() / np.abs(ref).max()
ok = err < 1e-12 and best == grid[int(np.argmin(ref))]
flag = torch.tensor([1 if ok else 0], device=d); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print(f"world={world} shard sizes ~{hi - lo}: total NLL (all-reduced) vs single-rank: rel err {err:.2e}; best s = {best}; "
          f"{'OK' if int(flag) else 'MISMATCH'}", flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
