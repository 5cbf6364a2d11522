"""Baselines this repo measures itself against: PyTorch DDP/NCCL + cuBLAS/cuDNN
renditions of the three reference models and an NCCL bandwidth sweep.  Nothing here is on the
product path."""
