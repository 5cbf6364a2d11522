"""Micro-benchmark of the persistent LSTM kernels: time per call, with/without saving."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from psb200.ops import cuda as K

def timeit(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n * 1e3

for B, T in [(128, 28), (128, 56), (128, 1), (16, 28)]:
    I, H = 28, 128
    x = torch.rand(B, T, I, device="cuda")
    kernel = (torch.rand(I + H, 4 * H, device="cuda") - 0.5) * 0.2
    W, Wt = K.BMat.from_tensor(kernel), K.BMat.from_tensor(kernel.t().contiguous())
    bias = torch.zeros(4 * H, device="cuda")
    h, ht = K.BMat(B, H), K.BMat(H, B)
    saved = K.LstmSaved(B, T, I, H)
    dh = torch.randn(B, H, device="cuda")
    db = torch.zeros(4 * H, device="cuda")
    t_inf = timeit(lambda: K.lstm_fwd(x, Wt, bias, 1.0, h, None, None))
    t_trn = timeit(lambda: K.lstm_fwd(x, Wt, bias, 1.0, h, ht, saved))
    t_bwd = timeit(lambda: K.lstm_bwd(W, dh, saved, db))
    print("B=%d T=%d  fwd(infer) %.1f us  fwd(train) %.1f us  bwd %.1f us" % (B, T, t_inf, t_trn, t_bwd))
