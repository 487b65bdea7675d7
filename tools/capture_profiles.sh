#!/bin/bash
# Round-2 ncu evidence (run under gpurun on ONE GPU; every ncu pass follows a plain run of the same command that exited 0).
# Outputs go to gpurun_out/; tools/summarize_ncu.py turns them into the text files kept under profiles/.
set -u
OUT=gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-alt --no-other-configs"
$B > $OUT/plain_c3.log 2>&1 || { echo "plain c3 run failed"; exit 1; }
# 1. every launch of one config-3 run with its device time
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file $OUT/r02_launches_c3.csv $B > $OUT/ncu1.log 2>&1
# 2. full captures: both recurrence kernels, the chunk GEMMs, the fused projection, the bandwidth-bound kernels
ncu --set full --clock-control none --import-source on -k regex:"rec_(fwd|bwd)_tc" -s 60 -c 4 -f -o $OUT/r02_rec_c3 $B > $OUT/ncu2.log 2>&1
ncu --set full --clock-control none -k regex:"gemm_bf16_tc" -s 200 -c 6 -f -o $OUT/r02_gemm_c3 $B > $OUT/ncu3.log 2>&1
ncu --set full --clock-control none -k regex:"proj_xent|embed_gather|embed_scatter|clip_adam|sqnorm|colsum_kernel" -s 12 -c 14 -f -o $OUT/r02_hbm_c3 $B > $OUT/ncu4.log 2>&1
# 3. one recurrence launch + one weight-gradient GEMM of configs 4 and 5
for c in c4 c5; do
  BC="python bench.py --config $c --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-alt --no-other-configs"
  $BC > $OUT/plain_$c.log 2>&1 || { echo "plain $c run failed"; continue; }
  ncu --set full --clock-control none -k regex:"rec_(fwd|bwd)_tc" -s 6 -c 2 -f -o $OUT/r02_rec_$c $BC > $OUT/ncu_$c.log 2>&1
  ncu --set full --clock-control none -k regex:"gemm_bf16_tc" -s 60 -c 3 -f -o $OUT/r02_gemm_$c $BC > $OUT/ncu_g$c.log 2>&1
done
# the reports are 30-40 MB each (gpurun brings back at most 64 MB): keep their raw pages as CSV, drop the reports
for f in $OUT/r02_*.ncu-rep; do
  ncu -i $f --page raw --csv > ${f%.ncu-rep}.raw.csv 2>/dev/null
  rm -f $f
done
ls -la $OUT/
