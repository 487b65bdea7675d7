cd $GRAFT_REPO_ROOT
OUT=gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.build(); g.smoke()" > $OUT/smoke2.log 2>&1 || { echo smoke failed; tail -5 $OUT/smoke2.log; exit 1; }
( time timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1000 --csv --log-file $OUT/smoke_launches2.csv python -c "import __graft_entry__ as g; g.build(); g.smoke()" > $OUT/smoke_ncu2.log 2>&1 ) 2>&1 | grep real
B="python bench.py --config c1 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-alt --no-other-configs"
$B > $OUT/plain_c1.log 2>&1 || { echo "plain c1 failed"; tail -5 $OUT/plain_c1.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"rec_(fwd|bwd)_tc" -s 40 -c 4 -f -o $OUT/r02_rec_c1 $B > $OUT/ncu_c1.log 2>&1
ncu -i $OUT/r02_rec_c1.ncu-rep --page raw --csv > $OUT/r02_rec_c1.raw.csv 2>/dev/null; rm -f $OUT/r02_rec_c1.ncu-rep
tail -3 $OUT/ncu_c1.log; ls -la $OUT | tail -5
