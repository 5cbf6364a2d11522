#!/usr/bin/env python
"""Headline benchmark: examples/sec of the distributed_lstm training step (BASELINE.json:
"examples/sec at 1/2/4/8 B200 (LSTM/GAN/AE)", config "distributed_lstm on synthetic MNIST
rows (28 steps x 28 features) sync mode"), one rank per GPU, weak scaling (128 examples per
GPU per step, lstm:29).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--model lstm|gan|encoder|scaled_lstm]
                    [--mode sync|async] [--impl psb200|reference|torch_ddp|torch_ddp_graph]

The product arm also runs the strongest baseline (PyTorch + cuBLAS/cuDNN + NCCL, the whole step
captured in one CUDA graph: ``baseline/torch_ddp.py``) in the SAME invocation -- same lease, same
clocks -- and reports it under ``baseline_same_run``; ``vs_baseline`` is computed from that run.

For N > 1 launch with ``python -m torch.distributed.run --nproc-per-node N ... bench.py
--gpus N ...``.  Rank 0 prints ONE JSON line.

Timing: every step is bracketed by its own pair of CUDA events on the launching stream;
between steps a 256 MiB buffer is overwritten to flush the 126 MB L2 (outside the event
pairs); the reported time is the sum of the K step durations, max over ranks.  The e2e arm
drives the public API: pinned-host batch -> H2D -> step -> D2H of loss and global_step, every
step, through ``DeviceTrainer.step_async`` (pipelined; headline) and ``DeviceTrainer.step``
(host blocks on every step; reported as ``blocking_api_*``).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PER_GPU_BATCH = {"lstm": 128, "gan": 128, "encoder": 256, "scaled_lstm": 64}
MODEL_LABEL = {"lstm": "distributed_lstm BasicLSTM(128) on 28x28 rows + linear-10, SGD 1e-3",
               "gan": "distributed_gan G 100-256-784 / D 784-256-1, 2x Adam 2e-4",
               "encoder": "distributed_encoder 784-256-128-256-784, RMSProp 0.01",
               "scaled_lstm": "scaled 2-layer LSTM hidden 4096 seq 512 bf16, SGD"}


from psb200.utils.clocks import ClockSampler  # noqa: E402


# ----------------------------------------------------------------------------- helpers
def _dist_setup(n_gpus: int):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != n_gpus:
        if world == 1 and n_gpus > 1:
            raise SystemExit("bench.py --gpus %d must be launched with torch.distributed.run "
                             "--nproc-per-node %d" % (n_gpus, n_gpus))
        n_gpus = world
    torch.cuda.set_device(local)
    group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # keep stdout to the one JSON line
        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local))
        group = dist.new_group(backend="gloo")
    return rank, world, local, group


def _max_over_ranks(x: float, world: int) -> float:
    if world == 1:
        return x
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _barrier(world: int):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def synthetic_batches(model: str, batch: int, n: int, seed: int, dims=None):
    """MNIST-shaped synthetic data in pinned host memory (images in [0,1], one-hot labels)."""
    import torch
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n):
        b = {}
        if model == "scaled_lstm":
            T, I = dims["timesteps"], dims["num_input"]
            b["x"] = torch.rand(T, batch, I, generator=g).pin_memory()
            b["y"] = torch.nn.functional.one_hot(torch.randint(0, 10, (batch,), generator=g), 10).float().pin_memory()
        elif model == "lstm":
            b["x"] = torch.rand(batch, 28, 28, generator=g).pin_memory()
            b["y"] = torch.nn.functional.one_hot(torch.randint(0, 10, (batch,), generator=g), 10).float().pin_memory()
        elif model == "gan":
            b["x"] = torch.rand(batch, 784, generator=g).pin_memory()
            b["z"] = (torch.rand(batch, 100, generator=g) * 2 - 1).pin_memory()
        else:
            b["x"] = torch.rand(batch, 784, generator=g).pin_memory()
        out.append(b)
    return out


def _measured_baseline(model: str, n_gpus: int):
    path = os.path.join(ROOT, "baseline", "measured.json")
    try:
        with open(path) as f:
            d = json.load(f)
        return d.get(model, {}).get(str(n_gpus), {}).get("examples_per_sec")
    except Exception:
        return None


# ----------------------------------------------------------------------------- timed loops
def timed_device_steps(step_fn, steps: int, world: int, flush_buf):
    """Sum of per-step CUDA-event durations with an L2 flush between steps; max over ranks."""
    import torch
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    _barrier(world)
    for i in range(steps):
        flush_buf.fill_(i & 0xFF)          # 256 MiB write: evicts the previous step's lines from L2
        starts[i].record()
        step_fn(i)
        ends[i].record()
    torch.cuda.synchronize()
    _barrier(world)
    ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    return _max_over_ranks(ms, world)


def timed_e2e_steps(step_fn, steps: int, world: int):
    """K public-API steps (H2D + step + D2H each) between two CUDA events; max over ranks."""
    import torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _barrier(world)
    e0.record()
    for i in range(steps):
        step_fn(i)
    e1.record()
    torch.cuda.synchronize()
    _barrier(world)
    return _max_over_ranks(e0.elapsed_time(e1), world)


# ----------------------------------------------------------------------------- arms
def run_psb200(args):
    import torch
    from psb200.ops import _lib
    from psb200.parallel.trainer import DeviceTrainer
    rank, world, local, group = _dist_setup(args.gpus)
    model = args.model
    B = args.batch or PER_GPU_BATCH[model]
    kw = {}
    if model == "gan":
        kw["device_noise"] = False      # z is fed from the host like np.random.uniform at gan:189
    dims = None
    if model == "scaled_lstm":
        dims = dict(num_input=args.hidden, timesteps=args.seq, num_hidden=args.hidden, num_layers=args.layers)
    tr = DeviceTrainer(model, rank, world, group, mode=args.mode, seed=args.seed, batch=B,
                       gan_update=args.gan_update, plan_kwargs=kw, model_kwargs=dims, bcast=bool(args.bcast),
                       partition=bool(args.partition), replicas_to_aggregate=args.replicas,
                       async_apply=args.async_apply)
    tr.initialize()
    batches = synthetic_batches(model, B, 2 if model == "scaled_lstm" else 8, args.seed * 1000 + rank, dims)
    dev_batches = [{k: v.cuda() for k, v in b.items()} for b in batches]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def dev_step(i):
        b = dev_batches[i % len(dev_batches)]
        for k, v in b.items():
            tr.plan.inputs[k].copy_(v, non_blocking=True)      # device->device, inputs stay resident
        tr.step_device()

    def e2e_step(i):
        tr.step(batches[i % len(batches)])

    def e2e_pipelined(i):
        # same work per step (H2D from pinned memory, step, D2H of loss + global_step); the host
        # does not wait for step i before submitting step i+1, results come back one call later
        tr.step_async(batches[i % len(batches)])
        if i == args.steps - 1:
            tr.drain()

    for i in range(args.warmup):
        dev_step(i)
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = _lib.launch_count()
    tr.store.push_profile(reset=True)
    ms = timed_device_steps(dev_step, args.steps, world, flush)
    push_prof = tr.store.push_profile(reset=True)
    launches = _lib.launch_count() - l0
    if tr.graph is not None:
        launches = tr.launches_per_step() * args.steps      # graph replays do not pass through the launcher
    for i in range(args.warmup):
        e2e_step(i)
    ms_e2e_sync = timed_e2e_steps(e2e_step, args.steps, world)
    for i in range(args.warmup):
        tr.step_async(batches[i % len(batches)])
    tr.drain()
    ms_e2e = timed_e2e_steps(e2e_pipelined, args.steps, world)
    clocks = sampler.stop() if rank == 0 else None
    used_graph = tr.graph is not None
    h2d_bytes, d2h_bytes = tr.h2d_bytes_per_step(batches[0]), tr.d2h_bytes_per_step()
    last = tr.step(batches[0])
    gstep = last["global_step"]
    tr.close()
    del tr, dev_batches, flush
    torch.cuda.empty_cache()
    # ---- the baseline arm, same process, same GPUs, right after the product arm
    base_run, base_err = None, None
    if args.baseline:
        try:
            from baseline.torch_ddp import run_graph
            base_run = run_graph(args, rank, world, local, sys.modules[__name__], keep_group=True)
        except Exception as e:      # the product number stands on its own; say why the ratio is missing
            base_err = "%s: %s" % (type(e).__name__, str(e)[:200])
    result = None
    if rank == 0:
        total = B * world * args.steps
        value = total / (ms / 1e3)
        e2e = total / (ms_e2e / 1e3)
        base = base_run["value"] if base_run else None
        base_label = "PyTorch cuBLAS/cuDNN + NCCL, whole step in one CUDA graph, measured in this run"
        if base is None:
            base = _measured_baseline(model, world)
            base_label = "EAGER PyTorch DDP rendition from baseline/measured.json (stored, other lease)" if base else None
        result = {
            "metric": "examples/sec", "value": round(value, 1), "unit": "examples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4),
            "higher_is_better": True, "scaling": "weak",
            "vs_baseline": round(value / base, 3) if base else None,
            "dtype": "bf16", "data": "synthetic MNIST-shaped, random-init weights",
            "impl": "psb200",
            "config": {"model": MODEL_LABEL[model] if dims is None else "scaled %d-layer LSTM hidden %d seq %d bf16, SGD" % (args.layers, args.hidden, args.seq),
                       "global_batch": B * world, "seq_len": 28 if model == "lstm" else (args.seq if dims else None),
                       "parallelism": "dp%d ps-sharded (%s%s%s%s)" % (
                           world, args.mode, "" if args.bcast else ", pull-fused",
                           ", R=%d of %d via ps service" % (args.replicas, world) if 0 < args.replicas < world else "",
                           ", owner-side apply via ps service" if args.mode == "async" and args.async_apply == "owner" and world > 1 else ""),
                       "gan_update": args.gan_update if model == "gan" else None,
                       "l2": "flushed between timed iterations (256 MiB write outside the event pairs)",
                       "cuda_graph": used_graph,
                       "baseline_for_vs": base_label},
            "clocks": clocks,
            "e2e": {"value": round(e2e, 1), "unit": "examples/s", "ms_per_step": round(ms_e2e / args.steps, 4),
                    "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": d2h_bytes,
                    "api": "DeviceTrainer.step_async: every step copies its batch from pinned host memory and "
                           "reads loss+global_step back; submission of step i+1 overlaps step i",
                    "blocking_api_value": round(total / (ms_e2e_sync / 1e3), 1),
                    "blocking_api_ms_per_step": round(ms_e2e_sync / args.steps, 4)},
            "gpu_launches": int(launches),
            "push_pull_us_per_step": {k: round(v, 2) for k, v in push_prof.items() if k != "calls"},
            "final": {"global_step": gstep, "losses": {k: v for k, v in last.items() if k != "global_step"}},
        }
        if base_run:
            result["baseline_same_run"] = {k: base_run[k] for k in ("impl", "value", "ms_per_step", "e2e", "clocks", "dtype")}
            result["e2e"]["vs_baseline_e2e"] = round(e2e / base_run["e2e"]["value"], 3)
        elif base_err:
            result["baseline_same_run"] = {"unavailable": base_err}
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return result


def run_reference(args):
    """The unmodified reference is three TensorFlow-1.11 scripts with no setup.py/pyproject
    (pip: "Directory '/root/reference' is not installable") and TensorFlow 1.x
    (tensorflow.contrib.rnn, tensorflow.examples.tutorials.mnist, tf.train.Supervisor) has no
    distribution for Python 3.12 in /opt/wheelhouse, so the arm cannot run here."""
    ok = False
    try:
        import tensorflow as tf  # noqa: F401
        ok = hasattr(tf, "app") and hasattr(tf, "contrib")
    except Exception:
        ok = False
    if not ok:
        return {"impl": "reference",
                "unavailable": "reference = TF-1.11 scripts (tf.contrib.rnn, tf.train.Supervisor); no "
                               "setup.py/pyproject to install and no TensorFlow 1.x wheel for py3.12 offline"}
    return {"impl": "reference", "unavailable": "TensorFlow 1.x found but MNIST data_dir is not shipped"}


def run_torch_ddp(args):
    from baseline.torch_ddp import run as run_ddp, run_graph
    rank, world, local, group = _dist_setup(args.gpus)
    fn = run_graph if args.impl == "torch_ddp_graph" else run_ddp
    return fn(args, rank, world, local, sys.modules[__name__])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--model", default="lstm", choices=["lstm", "gan", "encoder", "scaled_lstm"])
    ap.add_argument("--hidden", type=int, default=4096, help="scaled_lstm: hidden and input width")
    ap.add_argument("--seq", type=int, default=512, help="scaled_lstm: timesteps")
    ap.add_argument("--layers", type=int, default=2, help="scaled_lstm: stacked layers")
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (0 = the model's reference value)")
    ap.add_argument("--mode", default="sync", choices=["sync", "async"])
    ap.add_argument("--gan_update", default="simultaneous", choices=["simultaneous", "alternating"])
    ap.add_argument("--impl", default="psb200", choices=["psb200", "reference", "torch_ddp", "torch_ddp_graph"])
    ap.add_argument("--baseline", type=int, default=1, help="1: also run the CUDA-graphed PyTorch baseline in this invocation")
    ap.add_argument("--partition", type=int, default=1, help="1: big variables are split in row blocks over all ranks")
    ap.add_argument("--replicas", type=int, default=0, help="sync: aggregate the first R < N gradients (ps service thread)")
    ap.add_argument("--async_apply", default="worker", choices=["worker", "owner"],
                    help="async: who applies a gradient -- the pushing worker (remote RMW) or the owner's ps service")
    ap.add_argument("--bcast", type=int, default=1, help="1: push kernel refreshes every rank's arena; 0: consumers pull from the owner")
    ap.add_argument("--seed", type=int, default=0)
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    # stdout carries exactly ONE line, the JSON: anything libraries print there (NCCL's version
    # banner, warnings) is sent to stderr for the duration of the run
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(obj) + "\n").encode())

    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) == 0:
            emit(run_reference(args))
        return
    res = run_psb200(args) if args.impl == "psb200" else run_torch_ddp(args)
    if res is not None:
        emit(res)


if __name__ == "__main__":
    main()
