#!/bin/bash
# One multi-GPU box session: correctness of the fused push at N ranks, then the BASELINE.md rows.
#   bench/run_box.sh N OUTDIR [quick]
N=$1; OUT=$2; QUICK=$3
mkdir -p $OUT
P=29600
tr() { P=$((P+1)); python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P "$@"; }
t() { name=$1; shift; tr tests/multigpu_worker.py "$@" > $OUT/test_$name.log 2>&1; echo "test $name: $(grep -c MULTIGPU_OK $OUT/test_$name.log) ok"; }
b() { name=$1; shift; tr bench.py --gpus $N --steps 200 --warmup 20 "$@" > $OUT/$name.json 2> $OUT/$name.err || echo "FAILED $name"; python - $OUT/$name.json <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(sys.argv[1], d["value"], d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], d.get("push_pull_us_per_step"),
          "base", (d.get("baseline_same_run") or {}).get("ms_per_step"), "gs", d["final"]["global_step"])
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
}
t lstm_sync --model lstm --mode sync
t gan_sync --model gan --mode sync
t lstm_rofn --model lstm --mode sync --replicas $((N-1)) --slow_rank 1 --steps 40
t gan_async_owner --model gan --mode async --async_apply owner --steps 20
b lstm_sync --model lstm
b encoder_sync --model encoder
b gan_sync --model gan
b gan_async_worker --model gan --mode async --baseline 0
b gan_async_owner --model gan --mode async --async_apply owner --baseline 0
b gan_alt_sync --model gan --gan_update alternating --baseline 0
b gan_alt_async --model gan --gan_update alternating --mode async --baseline 0
if [ -z "$QUICK" ]; then
  b lstm_rofn --model lstm --replicas $((N-1)) --baseline 0
  tr bench/sweep_push_pull.py --max_mb 1024 --out $OUT/push_pull_sweep_${N}gpu.json > $OUT/sweep.log 2>&1; tail -12 $OUT/sweep.log
fi
