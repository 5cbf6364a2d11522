"""Smallest useful workload for compute-sanitizer (T4): one GEMM with every epilogue output, one
LSTM forward/backward, the loss kernels and a single-rank push.  Run as
    compute-sanitizer --tool memcheck python bench/sanitize_small.py
(one tool per run; see /opt/skills/guides/B200_PROFILING.md)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from psb200.ops import cuda as K  # noqa: E402
from psb200.parallel.trainer import DeviceTrainer  # noqa: E402

a = K.BMat.from_tensor(torch.randn(128, 100, device="cuda"))
b = K.BMat.from_tensor(torch.randn(40, 100, device="cuda"))
y = K.BMat.from_tensor(torch.rand(128, 40, device="cuda"))
out, out_t = K.BMat(128, 40), K.BMat(40, 128)
of = torch.zeros(128, 40, device="cuda")
cs = torch.zeros(40, device="cuda")
K.gemm(a, b, bias=cs, act=1, yprev=y, dact=1, out=out, out_t=out_t, out_f32=of, colsum=cs)
b.remote = True
K.gemm(a, b, out=out)
x = torch.rand(16, 3, 28, device="cuda")
kern = torch.randn(156, 512, device="cuda") * 0.1
W, Wt = K.BMat.from_tensor(kern), K.BMat.from_tensor(kern.t().contiguous())
h, ht, sv = K.BMat(16, 128), K.BMat(128, 16), K.LstmSaved(16, 3, 28, 128)
K.lstm_fwd(x, Wt, torch.zeros(512, device="cuda"), 1.0, h, ht, sv)
K.lstm_bwd(W, torch.randn(16, 128, device="cuda"), sv, torch.zeros(512, device="cuda"))
torch.cuda.synchronize()
tr = DeviceTrainer("gan", use_graph=False, seed=0)
tr.initialize()
tr.step_device()
torch.cuda.synchronize()
tr.close()
print("SANITIZE_WORKLOAD_OK")
