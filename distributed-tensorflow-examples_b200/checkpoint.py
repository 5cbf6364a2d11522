"""Name-keyed checkpoints with the reference's file and variable naming.

``tf.train.Saver()`` + ``Supervisor(logdir=model_dir, save_model_secs=60)`` (``encoder:138,
147-153``) write ``<model_dir>/model.ckpt-<global_step>.{index,data-00000-of-00001,meta}`` and
a ``checkpoint`` state file, keep the newest five, and save every global variable under its
graph name (``Variable``, ``Variable_1`` .., ``Variable/RMSProp`` ..,
``beta1_power`` .., ``rnn/basic_lstm_cell/kernel``).  File names, the state-file syntax and the
key set are reproduced.  Two bundle formats: ``"psb200"`` (default; JSON index + raw
little-endian data, self-describing and diff-able) and ``"tf"`` (TensorFlow's own V2 bundle:
string-table ``.index`` + raw ``.data`` shard, written natively by ``csrc/tf_bundle.cpp``).
``restore``/``list_variables`` recognise either.  ``.meta`` is always this framework's JSON
placement table: a TF ``MetaGraphDef`` describes a TF graph, which does not exist here.
"""
from __future__ import annotations

import json
import os
import re
from typing import Dict, List, Optional

import numpy as np
import torch

_DT = {"float32": torch.float32, "int32": torch.int32}
_NP = {"float32": np.float32, "int32": np.int32}


def _state_path(model_dir: str) -> str:
    return os.path.join(model_dir, "checkpoint")


def read_state(model_dir: str) -> List[str]:
    """All checkpoint prefixes listed in the ``checkpoint`` file, oldest first."""
    p = _state_path(model_dir)
    if not os.path.exists(p):
        return []
    out = []
    with open(p) as f:
        for line in f:
            m = re.match(r'\s*all_model_checkpoint_paths:\s*"(.*)"', line)
            if m:
                out.append(m.group(1))
    return out


def latest_checkpoint(model_dir: str) -> Optional[str]:
    """``tf.train.latest_checkpoint``: full prefix of the newest checkpoint, or None."""
    p = _state_path(model_dir)
    if not os.path.exists(p):
        return None
    with open(p) as f:
        for line in f:
            m = re.match(r'\s*model_checkpoint_path:\s*"(.*)"', line)
            if m:
                prefix = m.group(1)
                if not os.path.isabs(prefix):
                    prefix = os.path.join(model_dir, prefix)
                return prefix if os.path.exists(prefix + ".index") else None
    return None


class Saver:
    def __init__(self, max_to_keep: int = 5, bundle_format: str = "psb200"):
        if bundle_format not in ("psb200", "tf"):
            raise ValueError("bundle_format must be 'psb200' or 'tf'")
        self.max_to_keep = max_to_keep
        self.bundle_format = bundle_format

    def save(self, values: Dict[str, torch.Tensor], model_dir: str, global_step: int,
             meta: Optional[dict] = None, basename: str = "model.ckpt") -> str:
        os.makedirs(model_dir, exist_ok=True)
        name = "%s-%d" % (basename, int(global_step))
        prefix = os.path.join(model_dir, name)
        if self.bundle_format == "tf":
            from . import tf_bundle
            tf_bundle.write_bundle(prefix, values)
            with open(prefix + ".meta", "w") as f:
                json.dump(meta or {}, f)
            self._update_state(model_dir, name)
            return prefix
        index, offset = {}, 0
        tmp = prefix + ".data-00000-of-00001.tmp"
        with open(tmp, "wb") as f:
            for key in sorted(values):
                t = values[key].detach().cpu().contiguous()
                dt = str(t.dtype).replace("torch.", "")
                raw = t.numpy().astype(_NP[dt], copy=False).tobytes()
                f.write(raw)
                index[key] = {"dtype": dt, "shape": list(t.shape), "offset": offset, "nbytes": len(raw)}
                offset += len(raw)
        os.replace(tmp, prefix + ".data-00000-of-00001")
        with open(prefix + ".index.tmp", "w") as f:
            json.dump({"format": "psb200-bundle-v1", "global_step": int(global_step), "tensors": index}, f)
        os.replace(prefix + ".index.tmp", prefix + ".index")
        with open(prefix + ".meta", "w") as f:
            json.dump(meta or {}, f)
        self._update_state(model_dir, name)
        return prefix

    def _update_state(self, model_dir: str, name: str) -> None:
        kept = [p for p in read_state(model_dir) if p != name]
        kept.append(name)
        while len(kept) > self.max_to_keep:
            old = kept.pop(0)
            for ext in (".index", ".data-00000-of-00001", ".meta"):
                try:
                    os.remove(os.path.join(model_dir, old + ext))
                except OSError:
                    pass
        tmp = _state_path(model_dir) + ".tmp"
        with open(tmp, "w") as f:
            f.write('model_checkpoint_path: "%s"\n' % name)
            for p in kept:
                f.write('all_model_checkpoint_paths: "%s"\n' % p)
        os.replace(tmp, _state_path(model_dir))

    @staticmethod
    def restore(prefix: str) -> Dict[str, torch.Tensor]:
        from . import tf_bundle
        if tf_bundle.is_tf_bundle(prefix):
            return tf_bundle.read_bundle(prefix)
        with open(prefix + ".index") as f:
            idx = json.load(f)
        out = {}
        with open(prefix + ".data-00000-of-00001", "rb") as f:
            blob = f.read()
        for key, e in idx["tensors"].items():
            arr = np.frombuffer(blob, dtype=_NP[e["dtype"]], count=int(np.prod(e["shape"], dtype=np.int64)) if e["shape"] else 1,
                                offset=e["offset"]).reshape(e["shape"]).copy()
            out[key] = torch.from_numpy(arr)
        return out

    @staticmethod
    def list_variables(prefix: str) -> Dict[str, List[int]]:
        """``tf.train.list_variables``: name -> shape."""
        from . import tf_bundle
        if tf_bundle.is_tf_bundle(prefix):
            return tf_bundle.list_variables(prefix)
        with open(prefix + ".index") as f:
            idx = json.load(f)
        return {k: e["shape"] for k, e in idx["tensors"].items()}
