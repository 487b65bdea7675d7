#!/bin/bash
# usage: sweep_env.sh <config> <precision> "VAR=a VAR2=b" "VAR=c" ...   one bench run per quoted env set
cfg=$1; prec=$2; shift 2
for e in "$@"; do
  out=$(env $e timeout 90 python bench.py --config $cfg --precision $prec --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-alt 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],3), round(d['phases_ms']['rec_fwd'],3), round(d['phases_ms']['rec_bwd'],3))
except Exception as e: print('fail')")
  echo "$e -> $out"
done
