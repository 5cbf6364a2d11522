"""Model families of the reference: auto-encoder, GAN, LSTM classifier (+ the scaled
2-layer LSTM of BASELINE.json's comm-bound configuration)."""
from .base import Model, StepOutput, TrainOp
from .encoder import AutoEncoder
from .gan import GAN
from .lstm import LSTMClassifier
from .scaled_lstm import ScaledLSTM


def make_model(name: str, **kw) -> Model:
    if name in ("encoder", "autoencoder", "distributed_encoder"):
        return AutoEncoder()
    if name in ("gan", "distributed_gan"):
        return GAN()
    if name in ("lstm", "distributed_lstm"):
        return LSTMClassifier()
    if name in ("scaled_lstm",):
        return ScaledLSTM(**kw)
    raise ValueError("unknown model %r" % name)
