"""T0: variable auto-naming, slot naming and round-robin placement vs SURVEY §2.4."""
import pytest
import torch

from psb200.models import GAN, AutoEncoder, LSTMClassifier
from psb200.placement import VariableRegistry, replica_device_setter, random_normal


def _build(model_cls, ps):
    reg = VariableRegistry(replica_device_setter(ps_tasks=ps))
    m = model_cls().build(reg)
    return m, reg


def test_encoder_checkpoint_keys():
    m, reg = _build(AutoEncoder, 1)
    names = reg.names()
    assert names[:9] == ["Variable"] + ["Variable_%d" % i for i in range(1, 9)]
    assert reg["Variable"].shape == (784, 256) and reg["Variable_3"].shape == (256, 784)
    assert reg["Variable_4"].shape == (256,) and reg["Variable_7"].shape == (784,)
    assert reg["Variable_8"].dtype == torch.int32 and reg["Variable_8"].shape == ()
    for k in range(8):
        base = "Variable" if k == 0 else "Variable_%d" % k
        assert base + "/RMSProp" in reg and base + "/RMSProp_1" in reg
    assert len(reg) == 25


def test_gan_checkpoint_keys():
    m, reg = _build(GAN, 1)
    assert reg["Variable"].shape == (100, 256) and reg["Variable_3"].shape == (256, 1)
    assert reg["Variable_7"].shape == (1,)
    for n in ("beta1_power", "beta2_power", "beta1_power_1", "beta2_power_1"):
        assert n in reg and reg[n].shape == ()
    for k in range(8):
        base = "Variable" if k == 0 else "Variable_%d" % k
        assert base + "/Adam" in reg and base + "/Adam_1" in reg
    assert len(reg) == 29
    gen, disc = m.train_ops
    assert gen.var_list == ["Variable", "Variable_1", "Variable_4", "Variable_5"]
    assert disc.var_list == ["Variable_2", "Variable_3", "Variable_6", "Variable_7"]
    assert m.steps_per_run() == 2
    # beta powers follow the min-name variable of their var_list
    assert reg["beta1_power"].primary == "Variable" and reg["beta1_power_1"].primary == "Variable_2"


def test_lstm_checkpoint_keys():
    m, reg = _build(LSTMClassifier, 1)
    assert reg.names() == ["Variable", "Variable_1", "rnn/basic_lstm_cell/kernel",
                           "rnn/basic_lstm_cell/bias", "Variable_2"]
    assert reg["rnn/basic_lstm_cell/kernel"].shape == (156, 512)
    assert reg["rnn/basic_lstm_cell/bias"].shape == (512,)


@pytest.mark.parametrize("ps", [1, 2, 3, 8])
@pytest.mark.parametrize("cls", [AutoEncoder, GAN, LSTMClassifier])
def test_round_robin(cls, ps):
    m, reg = _build(cls, ps)
    primaries = [s for s in reg if s.kind in ("primary", "step")]
    for i, s in enumerate(primaries):
        assert s.ps_task == i % ps, (s.name, s.ps_task)
    for s in reg:
        if s.kind in ("slot", "nonslot"):
            assert s.ps_task == reg[s.primary].ps_task


def test_unique_names_with_explicit_collision():
    reg = VariableRegistry(replica_device_setter(ps_tasks=1))
    a = reg.variable(random_normal((2,)), (2,), name="Variable_1")
    b = reg.variable(random_normal((2,)), (2,))
    c = reg.variable(random_normal((2,)), (2,))
    assert (a.name, b.name) == ("Variable_1", "Variable")
    assert c.name == "Variable_2"       # Variable_1 is taken


def test_initial_values_shapes_and_determinism():
    m, reg = _build(AutoEncoder, 2)
    v1, v2 = reg.initial_values(seed=3), reg.initial_values(seed=3)
    for s in reg:
        assert tuple(v1[s.name].shape) == s.shape
        assert torch.equal(v1[s.name], v2[s.name])
    assert float(v1["Variable/RMSProp"].min()) == 1.0      # rms slot starts at one
    assert float(v1["Variable/RMSProp_1"].abs().max()) == 0.0
