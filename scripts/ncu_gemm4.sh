cd $GRAFT_REPO_ROOT
GEMM_ONLY="dec fc1,dec dfc2,dec proj,enc fc1" ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -c 8 -o gpurun_out/r01d_gemm4 python scripts/gemm_bench.py 2 > gpurun_out/r01d_gemm4.log 2>&1
tail -3 gpurun_out/r01d_gemm4.log
