"""Eager (un-graphed) training steps of one model for ncu: every kernel is a separate launch.

    python bench/profile_step.py --model lstm --steps 3
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from psb200.parallel.trainer import DeviceTrainer  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="lstm")
ap.add_argument("--steps", type=int, default=3)
a = ap.parse_args()
tr = DeviceTrainer(a.model, use_graph=False, seed=0)
tr.initialize()
for k, v in tr.plan.inputs.items():
    v.copy_(torch.rand_like(v))
if "y" in tr.plan.inputs:
    tr.plan.inputs["y"].zero_()
    tr.plan.inputs["y"][:, 3] = 1
for _ in range(a.steps):
    tr.step_device()
torch.cuda.synchronize()
print("done", tr.global_step())
tr.close()
