cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -x -q -m gpu -k "scatter or resume or gnorm or embedding" 2>&1 | tail -3
for c in c3 c1; do timeout 120 python tools/quick_bench.py $c bf16x3 20 2>&1 | grep -E "QB|phases|rror" | cut -c1-330; done
timeout 300 python tools/quick_bench.py c5 bf16x3 5 2>&1 | grep -E "phases|rror" | cut -c1-400
timeout 300 python tools/quick_bench.py c4 bf16 5 2>&1 | grep -E "phases|rror" | cut -c1-400
