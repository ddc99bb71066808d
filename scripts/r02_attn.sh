cd $GRAFT_REPO_ROOT
T=${1:-r02e}
for bo in 0 16 48 96; do echo "--- backoff $bo"; SSL4P_ATTN_BACKOFF=$bo timeout 300 python scripts/attn_bench.py 5 2>&1 | grep tcgen05; done
echo "--- PAIR=0 backoff 0"; SSL4P_ATTN_BACKOFF=0 SSL4P_ATTN_PAIR=0 timeout 300 python scripts/attn_bench.py 5 2>&1 | grep "T=50"| grep tcgen05
