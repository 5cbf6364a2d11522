"""Time the building blocks of the scaled LSTM step at full size (H=4096, T=512, B=64)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from psb200.ops import cuda as K

def timeit(fn, n=10, flush=None):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(n):
        if flush is not None: flush.fill_(1)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        tot += s.elapsed_time(e)
    return tot / n * 1e3

H, T, B = 4096, 512, 64
TB = T * B
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def mat(r, c): return K.BMat.from_tensor(torch.randn(r, c, device="cuda") * 0.05)
Wt = mat(4 * H, 2 * H)          # kernel^T [4H, I+H]
W = mat(2 * H, 4 * H)
X = mat(TB, H); XT = mat(H, TB)
h = mat(B, H)
dG = mat(TB, 4 * H); dGT = mat(4 * H, TB)
gx = torch.zeros(TB, 4 * H, device="cuda"); gh = torch.zeros(B, 4 * H, device="cuda")
gk = torch.zeros(2 * H, 4 * H, device="cuda")
dY = torch.zeros(TB, H, device="cuda"); dh = torch.zeros(B, H, device="cuda")
c = torch.zeros(B, H, device="cuda"); bias = torch.zeros(4 * H, device="cuda")
gates = torch.zeros(B, 4 * H, dtype=torch.bfloat16, device="cuda"); cs = torch.zeros(B, H, device="cuda")
hout = K.BMat(B, H)
wx_t, wh_t = Wt.cols_view(0, H), Wt.cols_view(H, 2 * H)
wx, wh = W.rows_view(0, H), W.rows_view(H, 2 * H)
fl = 2.0 * TB * H * 4 * H
for bn in (0, 128, 256):
    t = timeit(lambda: K.gemm(X, wx_t, out_f32=gx, block_n=bn), 3)
    print("big fwd GEMM [%d x %d x %d] bn=%d: %.0f us  %.0f TFLOP/s" % (TB, 4 * H, H, bn, t, fl / t / 1e6))
t = timeit(lambda: K.gemm(XT, dGT, out_f32=gk[:H]), 3)
print("wgrad GEMM [%d x %d x %d]: %.0f us  %.0f TFLOP/s" % (H, 4 * H, TB, t, fl / t / 1e6))
t = timeit(lambda: K.gemm(dG, wx, out_f32=dY), 3)
print("dY GEMM [%d x %d x %d]: %.0f us  %.0f TFLOP/s" % (TB, H, 4 * H, t, fl / t / 1e6))
for bn in (0, 32, 64, 128):
    t = timeit(lambda: K.gemm(h, wh_t, out_f32=gh, block_n=bn), 20, flush)
    print("step GEMM fwd [64 x %d x %d] bn=%d: %.1f us (%.0f GB/s of weights)" % (4 * H, H, bn, t, 4 * H * H * 2 / t / 1e3))
dg1 = dG.rows_view(0, B)
dhp = torch.zeros(8, B, H, device='cuda')
for bn, sk in ((64, 1), (64, 2), (64, 4), (128, 2), (128, 4)):
    t = timeit(lambda: K.gemm(dg1, wh, out_f32=dhp[0], block_n=bn, split_k=sk, split_stride=dhp.stride(0)), 20, flush)
    print("step GEMM bwd [64 x %d x %d] bn=%d split_k=%d: %.1f us" % (H, 4 * H, bn, sk, t))
t = timeit(lambda: K.lstm_cell_fwd(gx[:B], gh, bias, 1.0, c, hout, gates, cs), 20, flush)
print("cell fwd: %.1f us" % t)
dh2 = torch.zeros(2, B, H, device='cuda')
t = timeit(lambda: K.lstm_cell_bwd(dh2, None, gates, cs, cs, c, dg1, bias), 20, flush)
print("cell bwd: %.1f us" % t)
t = timeit(lambda: K.transpose_bf16(dG, dGT), 3)
print("transpose [%d x %d]: %.0f us" % (TB, 4 * H, t))
