"""M2 — the dense GAN of ``gan/distributed_gan.py``: G 100-256-784, D 784-256-1
(variables ``:122-133``, networks ``:76-93``, losses ``:142-143``, two
``AdamOptimizer(2e-4)`` with ``var_list`` and a shared ``global_step`` ``:146-159``).

One training call runs both ``train_gen`` and ``train_disc`` from a single forward pass
(``gan:193``): generator gradients flow through the discriminator into G's variables,
discriminator gradients come from both the real and the fake batch."""
from __future__ import annotations

from typing import Dict

import torch

from ..ops import reference as R
from ..optim import AdamOptimizer
from ..placement import VariableRegistry, glorot_normal_ref, zeros
from .base import Model, StepOutput

num_steps = 100000
batch_size = 128
learning_rate = 0.0002
image_dim = 784
gen_hidden_dim = 256
disc_hidden_dim = 256
noise_dim = 100


class GAN(Model):
    name = "gan"
    batch_size = batch_size
    num_steps = num_steps

    def build(self, registry: VariableRegistry) -> "GAN":
        self.registry = registry
        wshapes = {"gen_hidden1": (noise_dim, gen_hidden_dim), "gen_out": (gen_hidden_dim, image_dim),
                   "disc_hidden1": (image_dim, disc_hidden_dim), "disc_out": (disc_hidden_dim, 1)}
        bshapes = {"gen_hidden1": (gen_hidden_dim,), "gen_out": (image_dim,),
                   "disc_hidden1": (disc_hidden_dim,), "disc_out": (1,)}
        for k, s in wshapes.items():
            self.var["w_" + k] = registry.variable(glorot_normal_ref(s), s)
        for k, s in bshapes.items():
            self.var["b_" + k] = registry.variable(zeros(s), s)
        self.global_step = registry.global_step()
        v = self.var
        gen_vars = [v["w_gen_hidden1"], v["w_gen_out"], v["b_gen_hidden1"], v["b_gen_out"]]
        disc_vars = [v["w_disc_hidden1"], v["w_disc_out"], v["b_disc_hidden1"], v["b_disc_out"]]
        self._minimize("train_gen", AdamOptimizer(learning_rate), gen_vars, "gen_loss")
        self._minimize("train_disc", AdamOptimizer(learning_rate), disc_vars, "disc_loss")
        return self

    def step_reference(self, params: Dict[str, torch.Tensor], batch: Dict[str, torch.Tensor]) -> StepOutput:
        p = {k: params[s.name] for k, s in self.var.items()}
        x, z = batch["x"], batch["z"]
        B = x.shape[0]
        # generator
        gh = R.dense_fwd(z, p["w_gen_hidden1"], p["b_gen_hidden1"], R.ACT_RELU)
        gs = R.dense_fwd(gh, p["w_gen_out"], p["b_gen_out"], R.ACT_SIGMOID)
        # discriminator on [real; fake] with shared weights
        d_in = torch.cat([x, gs], dim=0)
        dh = R.dense_fwd(d_in, p["w_disc_hidden1"], p["b_disc_hidden1"], R.ACT_RELU)
        logit = R.dense_fwd(dh, p["w_disc_out"], p["b_disc_out"], R.ACT_NONE)
        gen_loss, disc_loss, g_fake_gen, g_real_disc, g_fake_disc = R.gan_losses(logit[:B], logit[B:])

        # ---- disc_loss -> discriminator variables (both halves of the batch)
        dlogit = torch.cat([g_real_disc, g_fake_disc], dim=0)
        ddh, dw_do, db_do = R.dense_bwd(dh, p["w_disc_out"], dlogit)
        dz1 = R.act_grad(dh, ddh, R.ACT_RELU)
        _, dw_d1, db_d1 = R.dense_bwd(d_in, p["w_disc_hidden1"], dz1, need_dx=False)
        disc_grads = {self.var["w_disc_hidden1"].name: dw_d1, self.var["w_disc_out"].name: dw_do,
                      self.var["b_disc_hidden1"].name: db_d1, self.var["b_disc_out"].name: db_do}

        # ---- gen_loss -> through D (fake half) -> generator variables
        dh_f = dh[B:]
        ddh_f = g_fake_gen @ p["w_disc_out"].t()
        dz1_f = R.act_grad(dh_f, ddh_f, R.ACT_RELU)
        dgs = dz1_f @ p["w_disc_hidden1"].t()
        dzo = R.act_grad(gs, dgs, R.ACT_SIGMOID)
        dgh, dw_go, db_go = R.dense_bwd(gh, p["w_gen_out"], dzo)
        dzh = R.act_grad(gh, dgh, R.ACT_RELU)
        _, dw_g1, db_g1 = R.dense_bwd(z, p["w_gen_hidden1"], dzh, need_dx=False)
        gen_grads = {self.var["w_gen_hidden1"].name: dw_g1, self.var["w_gen_out"].name: dw_go,
                     self.var["b_gen_hidden1"].name: db_g1, self.var["b_gen_out"].name: db_go}
        return StepOutput({"gen_loss": gen_loss, "disc_loss": disc_loss},
                          {"train_gen": gen_grads, "train_disc": disc_grads},
                          {"gen_sample": gs})
