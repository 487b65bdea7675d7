#!/bin/bash
# usage: tools/sweep_small.sh c3 "VAR=val VAR2=val" ...   (each arg = one env setting)
cfg=$1; shift
for e in "$@"; do
  out=$(env $e timeout 60 python bench.py --config $cfg --precision bf16 --steps 8 --warmup 3 --no-cpu-baseline --no-e2e --no-alt 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],3), round(d['phases_ms']['rec_fwd'],3), round(d['phases_ms']['rec_bwd'],3))
except Exception as e: print('fail')")
  echo "$e -> $out"
done
