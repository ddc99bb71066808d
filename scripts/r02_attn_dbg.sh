cd $GRAFT_REPO_ROOT
for d in 0 8 16 24 32 56 63 64 127 0; do echo "== SSL4P_ATTN_DEBUG=$d"; SSL4P_ATTN_DEBUG=$d timeout 120 python scripts/attn_bench.py 5 2>&1 | grep -E "T=197 H=16"; done
