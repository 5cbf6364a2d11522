"""List or print the variables of a checkpoint written by this framework (either bundle format).

    python -m psb200.tools.inspect_checkpoint --file_name=/tmp/checkpoints            # newest in the directory
    python -m psb200.tools.inspect_checkpoint --file_name=/tmp/checkpoints/model.ckpt-1200 --tensor_name=Variable_2
    python -m psb200.tools.inspect_checkpoint --file_name=... --all_tensors

The flags follow TensorFlow's ``inspect_checkpoint`` tool, the usual way to look inside what
``tf.train.Saver`` (``encoder:138``) wrote; the key set is the one of SURVEY §2.4.
"""
from __future__ import annotations

import argparse
import os
import sys
from typing import List, Optional

from ..checkpoint import Saver, latest_checkpoint
from .. import tf_bundle


def resolve(file_name: str) -> Optional[str]:
    """A directory means its newest checkpoint; otherwise ``file_name`` is the prefix itself."""
    if os.path.isdir(file_name):
        return latest_checkpoint(file_name)
    return file_name if os.path.exists(file_name + ".index") else None


def main(argv: Optional[List[str]] = None) -> int:
    ap = argparse.ArgumentParser(prog="inspect_checkpoint")
    ap.add_argument("--file_name", required=True, help="checkpoint prefix, or a model_dir (newest checkpoint)")
    ap.add_argument("--tensor_name", default="", help="print this tensor only")
    ap.add_argument("--all_tensors", action="store_true", help="print every tensor's values")
    a = ap.parse_args(argv)
    prefix = resolve(a.file_name)
    if prefix is None:
        print("no checkpoint found at %s" % a.file_name, file=sys.stderr)
        return 1
    fmt = "tf" if tf_bundle.is_tf_bundle(prefix) else "psb200"
    shapes = Saver.list_variables(prefix)
    values = Saver.restore(prefix)
    if a.tensor_name:
        if a.tensor_name not in shapes:
            print("tensor %r not in %s" % (a.tensor_name, prefix), file=sys.stderr)
            return 1
        print("tensor_name: ", a.tensor_name)
        print(values[a.tensor_name].numpy())
        return 0
    print("# %s (%s bundle, %d variables)" % (prefix, fmt, len(shapes)))
    for name in sorted(shapes):
        print("%s (DT_%s) %s" % (name, _dt(values[name]), shapes[name]))
        if a.all_tensors:
            print(values[name].numpy())
    return 0


def _dt(t) -> str:
    import torch
    return {torch.float32: "FLOAT", torch.int32: "INT32", torch.float64: "DOUBLE", torch.int64: "INT64",
            torch.bfloat16: "BFLOAT16", torch.float16: "HALF"}.get(t.dtype, str(t.dtype))


if __name__ == "__main__":
    sys.exit(main())
