#!/bin/bash
# usage: bench/run_scaling.sh N "models" "impls" [steps]   -> appends JSON lines to gpurun_out/scale_N.jsonl
N=$1; MODELS=${2:-"lstm gan encoder"}; IMPLS=${3:-"psb200 torch_ddp"}; STEPS=${4:-200}
mkdir -p gpurun_out
for m in $MODELS; do for impl in $IMPLS; do
  if [ "$N" = "1" ]; then
    timeout 300 python bench.py --gpus 1 --model $m --impl $impl --steps $STEPS --warmup 20 2>/dev/null | tail -1 >> gpurun_out/scale_$N.jsonl
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --model $m --impl $impl --steps $STEPS --warmup 20 2>/dev/null | grep '^{' | tail -1 >> gpurun_out/scale_$N.jsonl
  fi
done; done
cat gpurun_out/scale_$N.jsonl | cut -c1-200
