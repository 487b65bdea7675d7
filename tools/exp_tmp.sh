cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_bf16x3.py tests/test_gpu_tensor.py -x -q -m gpu 2>&1 | tail -4
timeout 900 python -m pytest tests/test_gpu_configs.py -x -q -m gpu -k "config4 or config5" -s 2>&1 | grep -E "passed|failed|worst|rror" | head
CRNN_VERBOSE=1 timeout 300 python tools/quick_bench.py c4 bf16x3,bf16 5 2>&1 | grep -E "QB|phases|rror|plan" | cut -c1-420
