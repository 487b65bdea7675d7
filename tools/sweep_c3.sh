#!/bin/bash
# sweep the recurrence decomposition on one config: prints ms/step and the rec phases
cfg=${1:-c3}
for hc in 16 32 48 64; do for nt in 2 4 8 13; do for sw in 0 1; do
  out=$(CRNN_TC_HC=$hc CRNN_TC_TILES=$nt CRNN_TC_SWAP=$sw timeout 60 python bench.py --config $cfg --precision bf16 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-alt 2>/dev/null | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],3), round(d['phases_ms']['rec_fwd'],3), round(d['phases_ms']['rec_bwd'],3))
except Exception as e: print('fail')")
  echo "hc=$hc tiles=$nt swap=$sw -> $out"
done; done; done
