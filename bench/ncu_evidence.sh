#!/bin/bash
# ncu evidence for every kernel of the library: one eager training step per model plus the full-size
# scaled-LSTM kernels.  The markdown summaries are produced on the GPU box (the reports themselves are
# too large to bring back).   bash bench/ncu_evidence.sh OUTDIR
OUT=${1:-gpurun_out/ncu_md}
mkdir -p $OUT
K='regex:gemm_bf16|lstm_|push_|_grad_kernel|cast_transpose|uniform_noise|transpose_bf16|wait_params'
one() {   # name  skip  count  title  command...
  name=$1; skip=$2; cnt=$3; title=$4; shift 4
  "$@" > $OUT/plain_$name.log 2>&1 || { echo "plain run of $name failed"; tail -3 $OUT/plain_$name.log; return; }
  ncu --set full --clock-control none -k "$K" -s $skip -c $cnt -o /tmp/prof_$name -f "$@" > $OUT/ncu_$name.log 2>&1
  python bench/ncu_summary.py /tmp/prof_$name.ncu-rep $OUT/ncu_$name.md "$title" > /dev/null 2>&1 && echo "$name: $(grep -c '^| `' $OUT/ncu_$name.md) kernels"
  tail -2 $OUT/ncu_$name.log > $OUT/ncu_$name.tail; rm -f $OUT/ncu_$name.log
}
# skip = the parameter refresh + the first step; count = one step
one lstm 5 4 "ncu: distributed_lstm training step (B=128, T=28, H=128): forward, BPTT+head, wgrad GEMM, fused push" python bench/profile_step.py --model lstm --steps 3
one encoder 15 14 "ncu: distributed_encoder training step (B=256): cast, 4 fwd GEMMs, MSE, 4 wgrad + 3 dgrad GEMMs, fused push" python bench/profile_step.py --model encoder --steps 3
one gan 17 16 "ncu: distributed_gan training step (B=128, simultaneous G/D update): cast, noise, 4 fwd GEMMs, loss, 8 backward GEMMs, fused push" python bench/profile_step.py --model gan --steps 3
one scaled 6 7 "ncu: scaled LSTM kernels at full size (H=4096, B=64): fused forward timestep, cell backward, split-K dh GEMM, batched GEMMs over 64 timesteps, transpose" python bench/ncu_scaled.py
du -sh $OUT
