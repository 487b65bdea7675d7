cd $GRAFT_REPO_ROOT
run3() { env "$@" timeout 120 python tools/quick_bench.py c3 $P 20 2>&1 | grep -E "QB|rror|phases" | cut -c1-420; }
P=bf16x3
run3 CRNN_X=0
run3 CRNN_TC_EXPT=64
P=bf16
run3 CRNN_X=0
run3 CRNN_TC_EXPT=64
