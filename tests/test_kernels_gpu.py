"""T2: LSTM forward/BPTT, loss and cast kernels vs the fp32 oracle (ops/reference.py)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return (a - b).abs().max().item() / max(1e-6, b.abs().max().item())


@pytest.mark.parametrize("B,T,I", [(128, 28, 28), (16, 5, 28), (48, 7, 12)])
def test_lstm_fwd_bwd(B, T, I):
    from psb200.ops import cuda as K_
    from psb200.ops import reference as R
    H = 128
    g = torch.Generator().manual_seed(0)
    x = torch.rand(B, T, I, generator=g).cuda()
    lim = (6.0 / (I + H + 4 * H)) ** 0.5
    kernel = ((torch.rand(I + H, 4 * H, generator=g) * 2 - 1) * lim * 2).cuda()
    bias = (torch.randn(4 * H, generator=g) * 0.1).cuda()
    dh_last = (torch.randn(B, H, generator=g) * 0.1).cuda()

    W = K_.BMat.from_tensor(kernel)              # [I+H, 4H]
    Wt = K_.BMat.from_tensor(kernel.t().contiguous())   # [4H, I+H]
    h_last, h_last_t = K_.BMat(B, H), K_.BMat(H, B)
    saved = K_.LstmSaved(B, T, I, H)
    K_.lstm_fwd(x, Wt, bias, 1.0, h_last, h_last_t, saved)
    torch.cuda.synchronize()

    kq = W.float()                                # bf16-rounded weights for the oracle
    h_ref, sv = R.lstm_fwd(x, kq, bias, 1.0)
    assert _rel(h_last.float(), h_ref) < 3e-2
    assert _rel(h_last_t.float().t(), h_ref) < 3e-2
    gates_ref = sv.gates.view(T, B, 4, H).permute(0, 2, 1, 3)
    assert _rel(saved.gates_tbh().float(), gates_ref) < 3e-2
    assert _rel(saved.c_tbh(), sv.c[1:]) < 3e-2
    xh_ref = sv.xh.permute(2, 0, 1).reshape(I + H, T * B)
    assert _rel(saved.xh_tb(), xh_ref) < 3e-2

    dbias = torch.zeros(4 * H, device="cuda")
    K_.lstm_bwd(W, dh_last, saved, dbias)
    dk = torch.zeros(I + H, 4 * H, device="cuda")
    K_.gemm(saved.xh_t, saved.dg_t, out_f32=dk)
    torch.cuda.synchronize()
    dk_ref, db_ref = R.lstm_bwd(dh_last, sv, kq)
    assert _rel(dbias, db_ref) < 5e-2
    assert _rel(dk, dk_ref) < 5e-2


@pytest.mark.parametrize("B,T,C", [(128, 28, 10), (40, 6, 16), (16, 3, 3)])
def test_lstm_bwd_fused_head(B, T, C):
    """psb_lstm_bwd_head == head GEMM + softmax-CE + head gradients + psb_lstm_bwd (fp32 oracle), twice in
    a row (the partial-sum scratch and its counter must reset themselves)."""
    from psb200.ops import cuda as K_
    from psb200.ops import reference as R
    H, I = 128, 28
    g = torch.Generator().manual_seed(1)
    x = torch.rand(B, T, I, generator=g).cuda()
    kernel = ((torch.rand(I + H, 4 * H, generator=g) * 2 - 1) * 0.2).cuda()
    bias = (torch.randn(4 * H, generator=g) * 0.1).cuda()
    w_out = torch.randn(H, C, generator=g).cuda()
    b_out = torch.randn(C, generator=g).cuda()
    labels = torch.eye(C)[torch.randint(0, C, (B,), generator=g)].cuda()
    W, Wt, Wo = K_.BMat.from_tensor(kernel), K_.BMat.from_tensor(kernel.t().contiguous()), K_.BMat.from_tensor(w_out)
    h_last, saved = K_.BMat(B, H), K_.LstmSaved(B, T, I, H)
    K_.lstm_fwd(x, Wt, bias, 1.0, h_last, None, saved)
    loss, correct = torch.zeros((), device="cuda"), torch.zeros((), device="cuda")
    logits = torch.zeros(B, C, device="cuda")
    # oracle on the kernel's own h_last (bf16) and bf16-rounded head weights
    hq, woq = h_last.float()[:, :H], Wo.float()[:, :C]
    z_ref = hq @ woq + b_out
    loss_ref, dz_ref = R.softmax_xent(z_ref, labels)
    _, sv = R.lstm_fwd(x, W.float(), bias, 1.0)
    dk_ref, db_ref = R.lstm_bwd(dz_ref @ woq.t(), sv, W.float())
    for accumulate in (True, False):
        if accumulate:
            dbias, dwo, dbo = torch.zeros(4 * H, device="cuda"), torch.zeros(H, C, device="cuda"), torch.zeros(C, device="cuda")
            dk = torch.zeros(I + H, 4 * H, device="cuda")
            K_.lstm_bwd_head(W, saved, dbias, h_last, Wo, b_out, labels, dwo, dbo, loss=loss, correct=correct, logits=logits)
        else:       # overwrite mode: stale contents must not survive, and the kernel clears dk itself
            dbias, dwo, dbo = torch.full((4 * H,), 7.0, device="cuda"), torch.full((H, C), 7.0, device="cuda"), torch.full((C,), 7.0, device="cuda")
            dk = torch.full((I + H, 4 * H), 7.0, device="cuda")
            K_.lstm_bwd_head(W, saved, dbias, h_last, Wo, b_out, labels, dwo, dbo, loss=loss, correct=correct,
                             logits=logits, accumulate=False, zero=dk)
        K_.gemm(saved.xh_t, saved.dg_t, out_f32=dk, split_k=2)
        torch.cuda.synchronize()
        assert _rel(logits, z_ref) < 1e-4
        assert abs(loss.item() - loss_ref.item()) < 1e-3 * max(1.0, abs(loss_ref.item()))
        assert correct.item() == (z_ref.argmax(1) == labels.argmax(1)).sum().item()
        assert _rel(dwo, hq.t() @ dz_ref) < 1e-3
        assert _rel(dbo, dz_ref.sum(0)) < 1e-3
        assert _rel(dbias, db_ref) < 5e-2
        assert _rel(dk, dk_ref) < 5e-2


def test_cast_transpose():
    from psb200.ops import cuda as K_
    x = torch.randn(130, 100, device="cuda")
    out, out_t = K_.BMat(130, 100), K_.BMat(100, 130)
    out.t.fill_(7)
    K_.cast_transpose(x, out, out_t)
    torch.cuda.synchronize()
    assert _rel(out.float(), x) < 1e-2 and _rel(out_t.float(), x.t()) < 1e-2
    assert out.t[:, 100:].abs().max().item() == 0


def test_mse_sigmoid_grad():
    from psb200.ops import cuda as K_
    from psb200.ops import reference as R
    M, N = 256, 784
    y = torch.rand(M, N, device="cuda")
    x = torch.rand(M, N, device="cuda")
    loss = torch.zeros((), device="cuda")
    dz, dz_t = K_.BMat(M, N), K_.BMat(N, M)
    cs = torch.zeros(N, device="cuda")
    K_.mse_sigmoid_grad(y, x, loss, dz, dz_t, cs)
    torch.cuda.synchronize()
    l_ref, dy = R.mse_loss(y, x)
    dz_ref = R.act_grad(y, dy, R.ACT_SIGMOID)
    assert abs(loss.item() - l_ref.item()) < 1e-5
    assert _rel(dz.float(), dz_ref) < 1e-2 and _rel(dz_t.float().t(), dz_ref) < 1e-2
    assert _rel(cs, dz_ref.sum(0)) < 1e-3


def test_gan_loss_grad():
    from psb200.ops import cuda as K_
    from psb200.ops import reference as R
    B = 128
    logit = torch.randn(2 * B, 1, device="cuda") * 2
    losses = torch.zeros(2, device="cuda")
    d_disc, d_disc_t, d_gen = K_.BMat(2 * B, 1), K_.BMat(1, 2 * B), K_.BMat(B, 1)
    db = torch.zeros(1, device="cuda")
    K_.gan_loss_grad(logit, B, losses, d_disc, d_disc_t, d_gen, db)
    torch.cuda.synchronize()
    gl, dl, g_fg, g_rd, g_fd = R.gan_losses(logit[:B], logit[B:])
    assert abs(losses[0].item() - gl.item()) < 1e-4 and abs(losses[1].item() - dl.item()) < 1e-4
    ref_disc = torch.cat([g_rd, g_fd], 0)
    assert _rel(d_disc.float(), ref_disc) < 1e-2 and _rel(d_disc_t.float().t(), ref_disc) < 1e-2
    assert _rel(d_gen.float(), g_fg) < 1e-2
    assert abs(db.item() - ref_disc.sum().item()) < 1e-4


def test_softmax_xent_grad():
    from psb200.ops import cuda as K_
    from psb200.ops import reference as R
    B, C = 128, 10
    logits = torch.randn(B, C, device="cuda") * 3
    labels = torch.nn.functional.one_hot(torch.randint(0, C, (B,)), C).float().cuda()
    loss = torch.zeros((), device="cuda")
    correct = torch.zeros((), device="cuda")
    dz, dz_t = K_.BMat(B, C), K_.BMat(C, B)
    cs = torch.zeros(C, device="cuda")
    K_.softmax_xent_grad(logits, labels, loss, dz, dz_t, cs, correct)
    torch.cuda.synchronize()
    l_ref, g_ref = R.softmax_xent(logits, labels)
    assert abs(loss.item() - l_ref.item()) < 1e-4
    assert _rel(dz.float(), g_ref) < 1e-2 and _rel(dz_t.float().t(), g_ref) < 1e-2
    assert _rel(cs, g_ref.sum(0)) < 1e-3
    assert correct.item() == (logits.argmax(1) == labels.argmax(1)).sum().item()
