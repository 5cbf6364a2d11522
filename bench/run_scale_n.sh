#!/bin/bash
# usage: scale.sh N OUT  -- three reference models + the scaled LSTM at N GPUs
N=$1; OUT=$2; mkdir -p $OUT; P=29700
run() { name=$1; shift; P=$((P+1)); if [ $N -gt 1 ]; then python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N "$@" > $OUT/$name.json 2> $OUT/$name.err; else python bench.py "$@" > $OUT/$name.json 2> $OUT/$name.err; fi; python - $OUT/$name.json <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(sys.argv[1], d["value"], d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], d.get("push_pull_us_per_step"), "base", (d.get("baseline_same_run") or {}).get("ms_per_step"), "vs", d.get("vs_baseline"))
except Exception as e:
    print(sys.argv[1], "unreadable", e)
PY
tail -2 $OUT/$name.err | cut -c1-300; }
if [ -z "$SCALED_ONLY" ]; then
run lstm --model lstm --steps 200 --warmup 20
run encoder --model encoder --steps 200 --warmup 20
run gan --model gan --steps 200 --warmup 20
fi
if [ -z "$NO_SCALED" ]; then run scaled_lstm --model scaled_lstm --steps 4 --warmup 3; fi
