cd $GRAFT_REPO_ROOT
T=${1:-r02m}
GEMM_CASES=cls GEMM_CG=1,2 SSL4P_GEMM_SCHED=1 timeout 300 python scripts/gemm_bench.py 5 > gpurun_out/${T}_gemm_cls_sched1.txt 2>&1; cat gpurun_out/${T}_gemm_cls_sched1.txt
GEMM_CASES=cls GEMM_CG=1,2 SSL4P_GEMM_SCHED=0 timeout 300 python scripts/gemm_bench.py 5 > gpurun_out/${T}_gemm_cls_sched0.txt 2>&1; cat gpurun_out/${T}_gemm_cls_sched0.txt | tail -3
GEMM_CG=1,2 SSL4P_GEMM_SCHED=0 timeout 300 python scripts/gemm_bench.py 5 > gpurun_out/${T}_gemm_mae_sched0.txt 2>&1; tail -2 gpurun_out/${T}_gemm_mae_sched0.txt
GEMM_CG=1,2 SSL4P_GEMM_SCHED=1 timeout 300 python scripts/gemm_bench.py 5 > gpurun_out/${T}_gemm_mae_sched1.txt 2>&1; tail -2 gpurun_out/${T}_gemm_mae_sched1.txt
