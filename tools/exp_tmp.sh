cd $GRAFT_REPO_ROOT
for S in 50 10; do
DBG_MODEL=lstm DBG_H=128 DBG_B=50 DBG_S=$S DBG_REF=11 timeout 120 python tools/dbg_rec_timeline.py bf16x3 2>&1 | grep -E "life|rror"
done
