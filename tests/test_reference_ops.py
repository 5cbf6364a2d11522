"""T0: the oracle's hand-written backward formulas vs torch.autograd, optimizer math vs
closed-form NumPy, BasicLSTMCell gate order / forget bias."""
import numpy as np
import pytest
import torch

from psb200.models import GAN, AutoEncoder, LSTMClassifier
from psb200.ops import reference as R
from psb200.optim import (AdamOptimizer, GradientDescentOptimizer, RMSPropOptimizer,
                          apply_reference)
from psb200.placement import VariableRegistry, replica_device_setter

torch.manual_seed(0)


def _model(cls, scale=0.1):
    reg = VariableRegistry(replica_device_setter(ps_tasks=1))
    m = cls().build(reg)
    vals = reg.initial_values(seed=1)
    params = {}
    for s in reg.trainable():
        params[s.name] = (vals[s.name] * scale).double().requires_grad_(True)
    return m, params


def _check(m, params, batch, loss_of):
    out = m.step_reference({k: v.detach() for k, v in params.items()}, batch)
    for op in m.train_ops:
        loss = loss_of(op.loss)
        grads = torch.autograd.grad(loss, [params[n] for n in op.var_list], retain_graph=True)
        for n, g in zip(op.var_list, grads):
            got = out.grads[op.name][n]
            assert torch.allclose(got, g, rtol=1e-8, atol=1e-10), (op.name, n, (got - g).abs().max())
    return out


def test_encoder_backward_matches_autograd():
    m, params = _model(AutoEncoder)
    x = torch.rand(16, 784, dtype=torch.float64)
    p = {k: params[s.name] for k, s in m.var.items()}
    h = x
    for wk, bk in (("encoder_h1", "encoder_b1"), ("encoder_h2", "encoder_b2"),
                   ("decoder_h1", "decoder_b1"), ("decoder_h2", "decoder_b2")):
        h = torch.sigmoid(h @ p[wk] + p[bk])
    loss = ((x - h) ** 2).mean()
    out = _check(m, params, {"x": x}, lambda _: loss)
    assert torch.allclose(out.losses["loss"], loss.detach())


def test_gan_backward_matches_autograd():
    m, params = _model(GAN, scale=1.0)
    B = 8
    x = torch.rand(B, 784, dtype=torch.float64)
    z = torch.rand(B, 100, dtype=torch.float64) * 2 - 1
    p = {k: params[s.name] for k, s in m.var.items()}

    def G(z):
        return torch.sigmoid(torch.relu(z @ p["w_gen_hidden1"] + p["b_gen_hidden1"]) @ p["w_gen_out"] + p["b_gen_out"])

    def D(x):
        return torch.sigmoid(torch.relu(x @ p["w_disc_hidden1"] + p["b_disc_hidden1"]) @ p["w_disc_out"] + p["b_disc_out"])

    fake = D(G(z))
    real = D(x)
    losses = {"gen_loss": -torch.log(fake).mean(),
              "disc_loss": -(torch.log(real) + torch.log(1.0 - fake)).mean()}
    out = _check(m, params, {"x": x, "z": z}, lambda k: losses[k])
    for k in losses:
        assert torch.allclose(out.losses[k], losses[k].detach(), rtol=1e-9)


def test_lstm_backward_matches_autograd_and_cell_math():
    m, params = _model(LSTMClassifier, scale=1.0)
    B = 4
    x = torch.rand(B, 28, 28, dtype=torch.float64)
    y = torch.nn.functional.one_hot(torch.randint(0, 10, (B,)), 10).double()
    p = {k: params[s.name] for k, s in m.var.items()}
    # independent formulation of BasicLSTMCell (gate order i, j, f, o; forget_bias at run time)
    h = torch.zeros(B, 128, dtype=torch.float64)
    c = torch.zeros(B, 128, dtype=torch.float64)
    for t in range(28):
        g = torch.cat([x[:, t], h], 1) @ p["kernel"] + p["bias"]
        i, j, f, o = g[:, :128], g[:, 128:256], g[:, 256:384], g[:, 384:]
        c = c * torch.sigmoid(f + 1.0) + torch.sigmoid(i) * torch.tanh(j)
        h = torch.tanh(c) * torch.sigmoid(o)
    logits = h @ p["out_w"] + p["out_b"]
    loss = torch.nn.functional.cross_entropy(logits, y.argmax(1))
    out = _check(m, params, {"x": x, "y": y}, lambda _: loss)
    assert torch.allclose(out.losses["loss"], loss.detach())
    assert torch.allclose(out.extras["logits"], logits.detach())


def test_sgd_rmsprop_adam_closed_form():
    g = torch.tensor([0.5, -2.0, 0.0])
    w0 = torch.tensor([1.0, 2.0, 3.0])
    # SGD
    w = w0.clone()
    apply_reference(GradientDescentOptimizer(0.1), w, g, [])
    assert np.allclose(w.numpy(), w0.numpy() - 0.1 * g.numpy())
    # RMSProp: rms starts at ONE
    opt = RMSPropOptimizer(0.01)
    w, ms, mom = w0.clone(), torch.ones(3), torch.zeros(3)
    apply_reference(opt, w, g, [ms, mom])
    ms_np = 0.9 * 1.0 + 0.1 * g.numpy() ** 2
    step = 0.01 * g.numpy() / np.sqrt(ms_np + 1e-10)
    assert np.allclose(ms.numpy(), ms_np) and np.allclose(mom.numpy(), step)
    assert np.allclose(w.numpy(), w0.numpy() - step)
    # Adam with bias correction from the *current* beta powers
    opt = AdamOptimizer(2e-4)
    w, m, v = w0.clone(), torch.zeros(3), torch.zeros(3)
    apply_reference(opt, w, g, [m, v], beta_powers=(0.9, 0.999))
    lr_t = 2e-4 * np.sqrt(1 - 0.999) / (1 - 0.9)
    m_np, v_np = 0.1 * g.numpy(), 0.001 * g.numpy() ** 2
    assert np.allclose(m.numpy(), m_np) and np.allclose(v.numpy(), v_np)
    assert np.allclose(w.numpy(), w0.numpy() - lr_t * m_np / (np.sqrt(v_np) + 1e-8), rtol=1e-6)


def test_losses_values():
    lr, lf = torch.randn(16, 1), torch.randn(16, 1)
    gl, dl, *_ = R.gan_losses(lr, lf)
    assert torch.allclose(gl, -torch.log(torch.sigmoid(lf)).mean(), atol=1e-6)
    assert torch.allclose(dl, -(torch.log(torch.sigmoid(lr)) + torch.log(1 - torch.sigmoid(lf))).mean(), atol=1e-6)
    logits = torch.randn(8, 10)
    y = torch.nn.functional.one_hot(torch.randint(0, 10, (8,)), 10).float()
    loss, _ = R.softmax_xent(logits, y)
    assert torch.allclose(loss, torch.nn.functional.cross_entropy(logits, y.argmax(1)), atol=1e-6)
