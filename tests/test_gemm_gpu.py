"""T2: tcgen05 GEMM + epilogue vs an fp32 torch oracle (reference shapes incl. the awkward
K=100/156, N=1/10 of gan:57,126 and lstm:31-34)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _mats(M, N, K, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    a = torch.randn(M, K, generator=g) * 0.5
    b = torch.randn(N, K, generator=g) * 0.5
    return a.cuda(), b.cuda()


@pytest.mark.parametrize("M,N,K,bn", [
    (128, 32, 64, 32), (128, 64, 128, 64), (256, 256, 784, 0), (256, 128, 256, 32),
    (128, 784, 256, 64), (128, 1, 256, 0), (128, 10, 128, 16), (128, 256, 100, 0),
    (784, 256, 256, 128), (100, 256, 128, 0), (300, 520, 456, 256), (256, 784, 256, 256),
    (1024, 2048, 1024, 0),
])
def test_gemm_plain(M, N, K, bn):
    from psb200.ops import cuda as K_
    a, b = _mats(M, N, K)
    A, B = K_.BMat.from_tensor(a), K_.BMat.from_tensor(b)
    out = K_.BMat(M, N)
    out_t = K_.BMat(N, M)
    out_f = torch.full((M, N), float("nan"), device="cuda")
    K_.gemm(A, B, out=out, out_t=out_t, out_f32=out_f, block_n=bn)
    torch.cuda.synchronize()
    ref = A.float() @ B.float().t()
    tol = 2e-2 * max(1.0, ref.abs().max().item())
    assert torch.allclose(out_f, ref, atol=1e-3 * max(1.0, ref.abs().max().item()), rtol=1e-4)
    assert (out.float() - ref).abs().max().item() < tol
    assert (out_t.float().t() - ref).abs().max().item() < tol
    # pad columns stay zero
    if out.ld > N:
        assert out.t[:, N:].abs().max().item() == 0


@pytest.mark.parametrize("act", [0, 1, 2])
def test_gemm_bias_act(act):
    from psb200.ops import cuda as K_
    M, N, K = 256, 200, 784
    a, b = _mats(M, N, K, 1)
    a *= 0.2
    bias = torch.randn(N, device="cuda")
    A, B = K_.BMat.from_tensor(a), K_.BMat.from_tensor(b)
    out = K_.BMat(M, N)
    K_.gemm(A, B, bias=bias, act=act, out=out)
    torch.cuda.synchronize()
    ref = A.float() @ B.float().t() + bias
    if act == 1:
        ref = torch.sigmoid(ref)
    elif act == 2:
        ref = torch.relu(ref)
    assert (out.float() - ref).abs().max().item() < 2e-2 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("dact", [1, 2])
def test_gemm_dact_colsum(dact):
    from psb200.ops import cuda as K_
    M, N, K = 256, 128, 256
    a, b = _mats(M, N, K, 2)
    y = torch.rand(M, N, device="cuda") if dact == 1 else torch.randn(M, N, device="cuda").relu()
    Y = K_.BMat.from_tensor(y)
    A, B = K_.BMat.from_tensor(a), K_.BMat.from_tensor(b)
    out = K_.BMat(M, N)
    out_t = K_.BMat(N, M)
    colsum = torch.zeros(N, device="cuda")
    K_.gemm(A, B, alpha=0.5, yprev=Y, dact=dact, out=out, out_t=out_t, colsum=colsum)
    torch.cuda.synchronize()
    yf = Y.float()
    d = yf * (1 - yf) if dact == 1 else (yf > 0).float()
    ref = 0.5 * (A.float() @ B.float().t()) * d
    scale = max(1.0, ref.abs().max().item())
    assert (out.float() - ref).abs().max().item() < 2e-2 * scale
    assert (out_t.float().t() - ref).abs().max().item() < 2e-2 * scale
    assert (colsum - ref.sum(0)).abs().max().item() < 1e-2 * max(1.0, ref.sum(0).abs().max().item())


@pytest.mark.parametrize("M,N,K,bn", [(256, 256, 784, 0), (128, 784, 256, 64), (128, 1, 256, 0),
                                      (128, 256, 100, 32), (300, 520, 456, 128), (256, 512, 1024, 256)])
def test_gemm_pull_variant_matches(M, N, K, bn):
    """psb_gemm_bf16_pull -- B by TMA over the (peer) address, or by the producer warp's ld.global into
    swizzled smem -- == the plain TMA kernel == oracle."""
    from psb200.ops import cuda as K_
    a, b = _mats(M, N, K, 3)
    bias = torch.randn(N, device="cuda")
    A, B = K_.BMat.from_tensor(a), K_.BMat.from_tensor(b)
    o1, o2, o3 = K_.BMat(M, N), K_.BMat(M, N), K_.BMat(M, N)
    K_.gemm(A, B, bias=bias, act=1, out=o1, block_n=bn)
    B.remote = True
    K_.gemm(A, B, bias=bias, act=1, out=o2, block_n=bn)
    K_.gemm(A, B, bias=bias, act=1, out=o3, block_n=bn, pull_mode=1)
    torch.cuda.synchronize()
    ref = torch.sigmoid(A.float() @ B.float().t() + bias)
    assert (o2.float() - ref).abs().max().item() < 2e-2
    assert torch.equal(o1.t, o2.t) and torch.equal(o1.t, o3.t)


@pytest.mark.parametrize("dim,nblk,first", [(0, 4, 2), (0, 1, 0), (1, 4, 1), (0, 8, 7)])
def test_gemm_with_readiness_flags(dim, nblk, first):
    """Tiles gated on per-row-block readiness flags (the pull fused into the GEMM): with every flag
    already at the expected step the result equals the ungated GEMM up to the order in which the
    k-blocks are accumulated (dim 0 visits them starting at `first`); a flag that never arrives ends
    in the timeout status instead of a hang."""
    from psb200.ops import cuda as K_
    from psb200.store.nvl import ReadyInfo
    M, N, K = 256, 512, 768
    a, b = _mats(M, N, K, 5)
    bias = torch.randn(N, device="cuda")
    A, B = K_.BMat.from_tensor(a), K_.BMat.from_tensor(b)
    o1, o2 = torch.zeros(M, N, device="cuda"), torch.zeros(M, N, device="cuda")
    K_.gemm(A, B, bias=bias, out_f32=o1)
    flags = torch.full((nblk + 1,), 7, dtype=torch.int32, device="cuda")
    step = torch.tensor([7], dtype=torch.int32, device="cuda")
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    rows = ((K if dim == 0 else N) + nblk - 1) // nblk
    B.ready = ReadyInfo(flags.data_ptr(), step.data_ptr(), rows, dim, nblk, first, status.data_ptr(), int(2e8))
    binfo = ReadyInfo(flags.data_ptr() + 4 * nblk, step.data_ptr(), 1, 0, 1, 0, status.data_ptr(), int(2e8))
    K_.gemm(A, B, bias=bias, bias_ready=binfo, out_f32=o2)
    torch.cuda.synchronize()
    assert int(status) == 0
    assert torch.allclose(o1, o2, rtol=1e-4, atol=1e-3 * o1.abs().max().item())
    if dim == 1 or first == 0:
        assert torch.equal(o1, o2)
    flags[nblk - 1] = 6                       # one row block never becomes ready
    K_.gemm(A, B, bias=bias, out_f32=o2)
    torch.cuda.synchronize()
    assert int(status) == 2                   # PSB_PUSH_ST_TIMEOUT_PARAMS, after ~0.2 s


@pytest.mark.parametrize("split,stride", [(2, 0), (4, 0), (2, 1), (4, 1)])
def test_gemm_split_k(split, stride):
    """Split-K: atomics into a zeroed fp32 output (stride 0) or separate partial buffers."""
    from psb200.ops import cuda as K_
    M, N, K = 64, 256, 2048
    a, b = _mats(M, N, K, 7)
    A, B = K_.BMat.from_tensor(a), K_.BMat.from_tensor(b)
    ref = A.float() @ B.float().t()
    if stride == 0:
        out = torch.zeros(M, N, device="cuda")
        K_.gemm(A, B, out_f32=out, split_k=split)
        got = out
    else:
        parts = torch.full((split, M, N), float("nan"), device="cuda")
        K_.gemm(A, B, out_f32=parts[0], split_k=split, split_stride=parts.stride(0))
        got = parts.sum(0)
    torch.cuda.synchronize()
    assert torch.allclose(got, ref, atol=2e-3 * ref.abs().max().item(), rtol=1e-4)
