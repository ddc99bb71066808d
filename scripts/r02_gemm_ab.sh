cd $GRAFT_REPO_ROOT
T=${1:-r02n}
timeout 400 python -m pytest tests/test_ops_gpu.py -m gpu -q -x -k "gemm" 2>&1 | tail -2
GEMM_CG=1,2 SSL4P_GEMM_SCHED=1 timeout 300 python scripts/gemm_bench.py 5 > gpurun_out/${T}_gemm_mae.txt 2>&1; tail -3 gpurun_out/${T}_gemm_mae.txt
GEMM_CASES=cls GEMM_CG=1,2 SSL4P_GEMM_SCHED=1 timeout 300 python scripts/gemm_bench.py 5 > gpurun_out/${T}_gemm_cls.txt 2>&1; tail -3 gpurun_out/${T}_gemm_cls.txt
GEMM_CASES=large GEMM_CG=1,2 SSL4P_GEMM_SCHED=1 timeout 300 python scripts/gemm_bench.py 5 > gpurun_out/${T}_gemm_large.txt 2>&1; tail -3 gpurun_out/${T}_gemm_large.txt
nvidia-smi --query-gpu=name,serial,uuid --format=csv | tail -1
