cd $GRAFT_REPO_ROOT
for d in 0 1 2 3 4 7 0; do echo "== SSL4P_ATTN_DEBUG=$d"; SSL4P_ATTN_DEBUG=$d timeout 120 python scripts/attn_bench.py 5 2>&1 | grep -E "T=197 H=16|T=50"; done
