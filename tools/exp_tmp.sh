cd $GRAFT_REPO_ROOT
run() { env "$@" CRNN_VERBOSE=1 timeout 100 python tools/quick_bench.py custom $P 10 2>&1 | grep -E "QB|rror|plan" | cut -c1-220; }
P=bf16x3
export QB_CFG=lstm,256,2,50,512,65; run CRNN_X=0; run CRNN_TC_SUB=0
export QB_CFG=gru,256,2,50,512,65; run CRNN_X=0; run CRNN_TC_SUB=0
export QB_CFG=rnn,256,2,50,1024,65; run CRNN_X=0; run CRNN_TC_SUB=0
