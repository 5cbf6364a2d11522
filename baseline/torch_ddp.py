"""PyTorch DDP/NCCL rendition of the reference models — the fallback baseline BASELINE.json
mandates because the TF-1.11 scripts cannot run on this image.

Same architectures, same per-GPU batch, same optimizers and hyper-parameters as
``encoder:45-51,137``, ``gan:50-57,146-159``, ``lstm:27-34,96``; idiomatic PyTorch: ``nn.Linear``
(cuBLAS), ``nn.LSTM`` (cuDNN), bf16 autocast, ``DistributedDataParallel`` gradient all-reduce
(NCCL), stock ``torch.optim`` optimizers.  Timed exactly like the product arm (``bench.py``).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class AE(nn.Module):
    def __init__(self):
        super().__init__()
        self.l = nn.ModuleList([nn.Linear(784, 256), nn.Linear(256, 128), nn.Linear(128, 256), nn.Linear(256, 784)])
        for m in self.l:
            nn.init.normal_(m.weight, std=1.0)
            nn.init.normal_(m.bias, std=1.0)

    def forward(self, x):
        h = x
        for m in self.l:
            h = torch.sigmoid(m(h))
        return ((x - h.float()) ** 2).mean()


class LSTMNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.rnn = nn.LSTM(28, 128, batch_first=True)
        self.head = nn.Linear(128, 10)

    def forward(self, x, y):
        out, _ = self.rnn(x)
        logits = self.head(out[:, -1])
        return F.cross_entropy(logits.float(), y.argmax(1))


class StackedLSTM(nn.Module):
    """cuDNN rendition of the scaled configuration: nn.LSTM(H, H, num_layers) + linear head."""

    def __init__(self, hidden=4096, layers=2):
        super().__init__()
        self.rnn = nn.LSTM(hidden, hidden, num_layers=layers)      # time-major input [T, B, I]
        self.head = nn.Linear(hidden, 10)

    def forward(self, x, y):
        out, _ = self.rnn(x)
        return F.cross_entropy(self.head(out[-1]).float(), y.argmax(1))


class G(nn.Module):
    def __init__(self):
        super().__init__()
        self.a, self.b = nn.Linear(100, 256), nn.Linear(256, 784)

    def forward(self, z):
        return torch.sigmoid(self.b(torch.relu(self.a(z))))


class D(nn.Module):
    def __init__(self):
        super().__init__()
        self.a, self.b = nn.Linear(784, 256), nn.Linear(256, 1)

    def forward(self, x):
        return self.b(torch.relu(self.a(x)))       # logits


def _wrap(m, world, local):
    m = m.cuda()
    if world > 1:
        from torch.nn.parallel import DistributedDataParallel as DDP
        m = DDP(m, device_ids=[local], gradient_as_bucket_view=True)
    return m


def build_step(model: str, world: int, local: int, dims=None):
    """Returns (step_fn(batch) -> loss tensor, input names)."""
    if model == "encoder":
        net = _wrap(AE(), world, local)
        opt = torch.optim.RMSprop(net.parameters(), lr=0.01, alpha=0.9, eps=1e-10)

        def step(b):
            opt.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = net(b["x"])
            loss.backward()
            opt.step()
            return loss
        return step
    if model == "scaled_lstm":
        net = _wrap(StackedLSTM(dims["num_hidden"], dims["num_layers"]).to(torch.bfloat16), world, local)
        opt = torch.optim.SGD(net.parameters(), lr=0.001)

        def step(b):
            opt.zero_grad(set_to_none=True)
            loss = net(b["x"].to(torch.bfloat16), b["y"])
            loss.backward()
            opt.step()
            return loss
        return step
    if model == "lstm":
        net = _wrap(LSTMNet(), world, local)
        opt = torch.optim.SGD(net.parameters(), lr=0.001)

        def step(b):
            opt.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = net(b["x"], b["y"])
            loss.backward()
            opt.step()
            return loss
        return step
    g, d = _wrap(G(), world, local), _wrap(D(), world, local)
    og = torch.optim.Adam(g.parameters(), lr=2e-4)
    od = torch.optim.Adam(d.parameters(), lr=2e-4)

    def step(b):
        # both train ops from one forward pass, like gan:193
        og.zero_grad(set_to_none=True)
        od.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            fake = g(b["z"])
            l_real = d(b["x"]).float()
            l_fake = d(fake).float()
            gen_loss = F.softplus(-l_fake).mean()
            disc_loss = (F.softplus(-l_real) + F.softplus(l_fake)).mean()
        g_params = [p for p in g.parameters()]
        d_params = [p for p in d.parameters()]
        gg = torch.autograd.grad(gen_loss, g_params, retain_graph=True)
        gd = torch.autograd.grad(disc_loss, d_params)
        for p, gr in zip(g_params, gg):
            p.grad = gr
        for p, gr in zip(d_params, gd):
            p.grad = gr
        if world > 1:
            import torch.distributed as dist
            flat = torch.cat([p.grad.reshape(-1) for p in g_params + d_params])
            dist.all_reduce(flat)
            flat /= world
            o = 0
            for p in g_params + d_params:
                p.grad.copy_(flat[o:o + p.numel()].view_as(p))
                o += p.numel()
        og.step()
        od.step()
        return gen_loss + disc_loss
    return step


def build_graph_step(model: str, world: int, local: int, dims, static_batch):
    """The strongest form of the baseline: the WHOLE step -- forward, backward, one flat NCCL
    all-reduce of the gradients, optimizer -- captured once in a CUDA graph and replayed, so it pays
    one launch per step like the product arm does.  cuBLAS/cuDNN kernels, bf16 autocast, capturable
    torch.optim optimizers.  Returns (replay_fn, static loss tensor)."""
    import torch.distributed as dist
    if model == "encoder":
        nets = [AE().cuda()]
        opts = [torch.optim.RMSprop(nets[0].parameters(), lr=0.01, alpha=0.9, eps=1e-10, capturable=True)]

        def fwd_bwd(b):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = nets[0](b["x"])
            loss.backward()
            return loss
    elif model == "lstm":
        nets = [LSTMNet().cuda()]
        opts = [torch.optim.SGD(nets[0].parameters(), lr=0.001)]

        def fwd_bwd(b):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = nets[0](b["x"], b["y"])
            loss.backward()
            return loss
    elif model == "scaled_lstm":
        nets = [StackedLSTM(dims["num_hidden"], dims["num_layers"]).to(torch.bfloat16).cuda()]
        opts = [torch.optim.SGD(nets[0].parameters(), lr=0.001)]

        def fwd_bwd(b):
            loss = nets[0](b["x"].to(torch.bfloat16), b["y"])
            loss.backward()
            return loss
    else:
        g, d = G().cuda(), D().cuda()
        nets = [g, d]
        opts = [torch.optim.Adam(g.parameters(), lr=2e-4, capturable=True),
                torch.optim.Adam(d.parameters(), lr=2e-4, capturable=True)]
        g_params, d_params = list(g.parameters()), list(d.parameters())

        def fwd_bwd(b):
            with torch.autocast("cuda", dtype=torch.bfloat16):
                fake = g(b["z"])
                l_real = d(b["x"]).float()
                l_fake = d(fake).float()
                gen_loss = F.softplus(-l_fake).mean()
                disc_loss = (F.softplus(-l_real) + F.softplus(l_fake)).mean()
            gen_loss.backward(inputs=g_params, retain_graph=True)
            disc_loss.backward(inputs=d_params)
            return gen_loss + disc_loss
    params = [p for n in nets for p in n.parameters()]
    # gradients live in ONE flat buffer (params' .grad are views): one all-reduce, no bucket copies
    flat = torch.zeros(sum(p.numel() for p in params), device="cuda", dtype=params[0].dtype)
    o = 0
    for p in params:
        p.grad = flat[o:o + p.numel()].view_as(p)
        o += p.numel()

    def body():
        flat.zero_()
        loss = fwd_bwd(static_batch)
        if world > 1:
            dist.all_reduce(flat, op=dist.ReduceOp.AVG)
        for opt in opts:
            opt.step()
        return loss

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            body()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        static_loss = body()
    torch.cuda.synchronize()
    return graph.replay, static_loss


def run_graph(args, rank, world, local, bench, keep_group: bool = False):
    """``--impl torch_ddp_graph``: the CUDA-graphed baseline, timed exactly like the product arm."""
    model = args.model
    B = args.batch or bench.PER_GPU_BATCH[model]
    dims = None
    if model == "scaled_lstm":
        dims = dict(num_input=args.hidden, timesteps=args.seq, num_hidden=args.hidden, num_layers=args.layers)
    batches = bench.synthetic_batches(model, B, 2 if dims else 8, args.seed * 1000 + rank, dims)
    dev_batches = [{k: v.cuda() for k, v in b.items()} for b in batches]
    static = {k: v.clone() for k, v in dev_batches[0].items()}
    replay, static_loss = build_graph_step(model, world, local, dims, static)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    host_loss = torch.zeros(1).pin_memory()

    def dev_step(i):
        for k, v in dev_batches[i % len(dev_batches)].items():
            static[k].copy_(v, non_blocking=True)
        replay()

    def e2e_step(i):
        for k, v in batches[i % len(batches)].items():
            static[k].copy_(v, non_blocking=True)               # pinned host -> device
        replay()
        host_loss.copy_(static_loss.detach().reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for i in range(args.warmup):
        dev_step(i)
    torch.cuda.synchronize()
    sampler = bench.ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = bench.timed_device_steps(dev_step, args.steps, world, flush)
    for i in range(args.warmup):
        e2e_step(i)
    ms_e2e = bench.timed_e2e_steps(e2e_step, args.steps, world)
    clocks = sampler.stop() if rank == 0 else None
    res = None
    if rank == 0:
        total = B * world * args.steps
        res = {"metric": "examples/sec", "value": round(total / (ms / 1e3), 1), "unit": "examples/s",
               "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "bf16 autocast", "data": "synthetic MNIST-shaped, random-init weights",
               "impl": "torch_ddp_graph",
               "config": {"model": bench.MODEL_LABEL[model], "global_batch": B * world,
                          "parallelism": "dp%d, whole step in one CUDA graph, flat NCCL all-reduce" % world,
                          "l2": "flushed between timed iterations (256 MiB write outside the event pairs)",
                          "cuda_graph": True},
               "clocks": clocks,
               "e2e": {"value": round(total / (ms_e2e / 1e3), 1), "unit": "examples/s",
                       "ms_per_step": round(ms_e2e / args.steps, 4),
                       "h2d_bytes_per_step": sum(v.numel() * 4 for v in batches[0].values()),
                       "d2h_bytes_per_step": 4},
               "gpu_launches": 0, "final_loss": float(static_loss.detach())}
    if world > 1 and not keep_group:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return res


def run(args, rank, world, local, bench):
    model = args.model
    B = args.batch or bench.PER_GPU_BATCH[model]
    dims = None
    if model == "scaled_lstm":
        dims = dict(num_input=args.hidden, timesteps=args.seq, num_hidden=args.hidden, num_layers=args.layers)
    step = build_step(model, world, local, dims)
    batches = bench.synthetic_batches(model, B, 2 if dims else 8, args.seed * 1000 + rank, dims)
    dev_batches = [{k: v.cuda() for k, v in b.items()} for b in batches]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    host_loss = torch.zeros(1).pin_memory()

    def dev_step(i):
        step(dev_batches[i % len(dev_batches)])

    def e2e_step(i):
        b = {k: v.cuda(non_blocking=True) for k, v in batches[i % len(batches)].items()}
        loss = step(b)
        host_loss.copy_(loss.detach().reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for i in range(args.warmup):
        dev_step(i)
    torch.cuda.synchronize()
    sampler = bench.ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = bench.timed_device_steps(dev_step, args.steps, world, flush)
    for i in range(args.warmup):
        e2e_step(i)
    ms_e2e = bench.timed_e2e_steps(e2e_step, args.steps, world)
    clocks = sampler.stop() if rank == 0 else None
    res = None
    if rank == 0:
        total = B * world * args.steps
        res = {"metric": "examples/sec", "value": round(total / (ms / 1e3), 1), "unit": "examples/s",
               "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "bf16 autocast", "data": "synthetic MNIST-shaped, random-init weights",
               "impl": "torch_ddp",
               "config": {"model": bench.MODEL_LABEL[model], "global_batch": B * world,
                          "parallelism": "dp%d DDP/NCCL" % world,
                          "l2": "flushed between timed iterations (256 MiB write outside the event pairs)"},
               "clocks": clocks,
               "e2e": {"value": round(total / (ms_e2e / 1e3), 1), "unit": "examples/s",
                       "ms_per_step": round(ms_e2e / args.steps, 4),
                       "h2d_bytes_per_step": sum(v.numel() * 4 for v in batches[0].values()),
                       "d2h_bytes_per_step": 4},
               "gpu_launches": 0}
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    return res
