"""Gradient push / parameter pull bandwidth sweep.

For message sizes 4 KB .. 1 GB (fp32 parameters, world equal shards, one per rank):
  psb200 : ONE psb_push_apply launch = reduce-scatter (P2P loads) + SGD on the owner +
           all-gather (P2P stores of the refreshed copy into every rank's arena)
  nccl   : dist.reduce_scatter_tensor + torch SGD on the shard + dist.all_gather_into_tensor
Device-timed with CUDA events, max over ranks, L2 flushed between iterations for sizes that fit
in L2.  "busbw" uses the NCCL convention for RS+AG: 2 * (W-1)/W * bytes / time.

    torchrun --nproc-per-node N bench/sweep_push_pull.py [--max_mb 1024]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from psb200.models.base import Model  # noqa: E402
from psb200.optim import GradientDescentOptimizer  # noqa: E402
from psb200.placement import VariableRegistry, replica_device_setter, zeros  # noqa: E402
from psb200.store.nvl import NvlStore  # noqa: E402


class BlobModel(Model):
    """`world` equal fp32 vectors, one owned by each rank (round-robin placement)."""
    name = "blob"

    def __init__(self, numel_per_rank: int, world: int):
        super().__init__()
        self.n, self.world = numel_per_rank, world

    def build(self, registry):
        self.registry = registry
        for r in range(self.world):
            self.var["v%d" % r] = registry.variable(zeros((self.n,)), (self.n,))
        self.global_step = registry.global_step()
        self._minimize("train", GradientDescentOptimizer(0.01), list(self.var.values()), "loss")
        return self


def timed(fn, iters, flush):
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.fill_(1)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier()
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    t = torch.tensor([ts[len(ts) // 2]], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max_mb", type=int, default=1024)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    group = dist.new_group(backend="gloo")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    sizes = [4 << 10]
    while sizes[-1] < (a.max_mb << 20):
        sizes.append(sizes[-1] * 4)
    rows = []
    for nbytes in sizes:
        n_total = nbytes // 4
        n_rank = max(4, (n_total // world + 3) // 4 * 4)
        iters = 20 if nbytes <= (64 << 20) else 6
        fl = flush if nbytes <= (128 << 20) else None
        # ---- psb200 fused push
        reg = VariableRegistry(replica_device_setter(ps_tasks=world))
        model = BlobModel(n_rank, world).build(reg)
        st = NvlStore(model, reg, rank, world, group, mode="sync", bcast=True)
        st.assign({s.name: torch.zeros(s.shape, dtype=s.dtype) for s in reg})
        st.grads.fill_(1.0)
        dist.barrier()
        st.refresh_copies()
        for _ in range(3):
            st.push()
        torch.cuda.synchronize()
        t_psb = timed(st.push, iters, fl)
        prof = st.push_profile()
        ok = abs(float(st.vec("Variable").flatten()[0]) + 0.01 * (3 + iters)) < 1e-3
        st.close()
        # ---- NCCL baseline
        full = torch.ones(n_rank * world, device="cuda")
        shard_g = torch.empty(n_rank, device="cuda")
        shard_w = torch.zeros(n_rank, device="cuda")
        gathered = torch.empty(n_rank * world, device="cuda")

        def nccl_step():
            dist.reduce_scatter_tensor(shard_g, full, op=dist.ReduceOp.AVG)
            shard_w.add_(shard_g, alpha=-0.01)
            dist.all_gather_into_tensor(gathered, shard_w)

        for _ in range(3):
            nccl_step()
        torch.cuda.synchronize()
        t_nccl = timed(nccl_step, iters, fl)
        moved = 2.0 * (world - 1) / world * n_rank * world * 4
        row = {"bytes": n_rank * world * 4, "world": world, "psb200_us": round(t_psb * 1e3, 2),
               "nccl_us": round(t_nccl * 1e3, 2), "psb200_busbw_gbs": round(moved / (t_psb * 1e-3) / 1e9, 2),
               "nccl_busbw_gbs": round(moved / (t_nccl * 1e-3) / 1e9, 2), "speedup": round(t_nccl / t_psb, 2),
               "correct": ok, "apply_us": round(prof["apply_us"], 2)}
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    if rank == 0 and a.out:
        with open(a.out, "w") as f:
            json.dump(rows, f, indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
