# usage: bash scripts/r02_gemm.sh TAG : CTA-pair GEMM on one B200: tests (guarded by a timeout), then the per-shape table for both groups
cd $GRAFT_REPO_ROOT
T=${1:-r02b}
timeout 600 python -m pytest tests/test_ops_gpu.py -m gpu -q -x -k "gemm" > gpurun_out/${T}_gemm_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_gemm_pytest.log
tail -15 gpurun_out/${T}_gemm_pytest.log
GEMM_CG=1,2 timeout 600 python scripts/gemm_bench.py 5 > gpurun_out/${T}_gemm_bench.txt 2>&1
cat gpurun_out/${T}_gemm_bench.txt | tail -60
