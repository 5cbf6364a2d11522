"""The kernels of one scaled-LSTM timestep at full size (H=4096, B=64) plus one batched GEMM of
each kind, a few launches each, for `ncu --set full` (profiles/ncu_scaled.md):

    python bench/ncu_scaled.py [T]        # T = timesteps the batched GEMMs cover (default 64)
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from psb200.ops import cuda as K  # noqa: E402

H, B = 4096, 64
T = int(sys.argv[1]) if len(sys.argv) > 1 else 64
TB = T * B


def mat(r, c):
    return K.BMat.from_tensor(torch.randn(r, c, device="cuda") * 0.05)


Wt, W = mat(4 * H, 2 * H), mat(2 * H, 4 * H)          # kernel^T [4H, I+H], kernel [I+H, 4H]
wx_t, wh_t = Wt.cols_view(0, H), Wt.cols_view(H, 2 * H)
wx, wh = W.rows_view(0, H), W.rows_view(H, 2 * H)
X, XT = mat(TB, H), mat(H, TB)
dG, dGT = mat(TB, 4 * H), mat(4 * H, TB)
h, hout = mat(B, H), K.BMat(B, H)
gx = torch.zeros(TB, 4 * H, device="cuda")
gk = torch.zeros(2 * H, 4 * H, device="cuda")
dY = torch.zeros(TB, H, device="cuda")
c, cs = torch.zeros(B, H, device="cuda"), torch.zeros(B, H, device="cuda")
bias = torch.zeros(4 * H, device="cuda")
gates = torch.zeros(B, 4 * H, dtype=torch.bfloat16, device="cuda")
dhp = torch.zeros(2, B, H, device="cuda")
dg1 = dG.rows_view(0, B)
for _ in range(3):
    # forward timestep: recurrent GEMM + cell pointwise in one kernel
    K.lstm_step_fwd(h, wh_t, gx[:B], bias, 1.0, c, hout, gates, cs)
    # backward timestep: cell backward, then dh = dG Wh^T (split-K 2, partial buffers)
    K.lstm_cell_bwd(dhp, None, gates, cs, cs, c, dg1, bias)
    K.gemm(dg1, wh, out_f32=dhp[0], split_k=2, split_stride=dhp.stride(0), block_n=64, b_static=True)
# batched GEMMs of a layer (input projection, weight gradient, gradient to the layer below)
K.gemm(X, wx_t, out_f32=gx, b_static=True)
K.gemm(XT, dGT, out_f32=gk[:H])
K.gemm(dG, wx, out_f32=dY, b_static=True)
K.transpose_bf16(dG, dGT)
torch.cuda.synchronize()
print("NCU_SCALED_OK")
