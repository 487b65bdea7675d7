cd $GRAFT_REPO_ROOT
OUT=gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-alt --no-other-configs"
$B > $OUT/plain_c3c.log 2>&1 || { echo "plain c3 run failed"; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file $OUT/r02_launches_c3.csv $B > $OUT/ncu1c.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:"proj_xent|embed_gather|es_sum|es_finish|es_rank|es_place|es_offsets|clip_adam|sqnorm|colsum_kernel" -s 12 -c 18 -f -o $OUT/r02_hbm_c3 $B > $OUT/ncu4c.log 2>&1
ncu -i $OUT/r02_hbm_c3.ncu-rep --page raw --csv > $OUT/r02_hbm_c3.raw.csv 2>/dev/null; rm -f $OUT/r02_hbm_c3.ncu-rep
ls -la $OUT/r02_launches_c3.csv $OUT/r02_hbm_c3.raw.csv
