cd $GRAFT_REPO_ROOT
for d in 0 1 2 3 4 5 7; do echo "== SSL4P_GEMM_DEBUG=$d"; SSL4P_GEMM_DEBUG=$d GEMM_ONLY="dec fc1,dec dfc2,dec qkv,dec proj,enc fc1" python scripts/gemm_bench.py 3 2>&1 | grep -v "^sum"; done
