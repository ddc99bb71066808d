cd $GRAFT_REPO_ROOT
T=${1:-r02o}
timeout 400 python -m pytest tests/test_ops_gpu.py -m gpu -q -x -k "gemm" 2>&1 | tail -2
runc() { n=$1; shift
  env "$@" timeout 300 python scripts/bench_classifier.py --steps 30 --warmup 5 > gpurun_out/${T}_cls_$n.json 2> gpurun_out/${T}_cls_$n.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_cls_$n.json').read().strip().splitlines()[-1])
print('cls $n', round(d['value']), round(d['ms_per_step'],3))
PY
}
runb() { n=$1; shift
  env "$@" timeout 300 python bench.py --steps 30 --warmup 5 --extra none --no-cpu-baseline --no-e2e > gpurun_out/${T}_mae_$n.json 2> gpurun_out/${T}_mae_$n.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_mae_$n.json').read().strip().splitlines()[-1])
print('mae $n', round(d['value']), round(d['ms_per_step'],3), d['clocks']['sm_mhz'])
PY
}
for rep in a b; do
runc sched0_$rep SSL4P_GEMM_SCHED=0
runc sched1_$rep SSL4P_GEMM_SCHED=1
runc sched1_cg1_$rep SSL4P_GEMM_SCHED=1 SSL4P_GEMM_CG=1
runb sched0_$rep SSL4P_GEMM_SCHED=0
runb sched1_$rep SSL4P_GEMM_SCHED=1
runb sched1_cg1_$rep SSL4P_GEMM_SCHED=1 SSL4P_GEMM_CG=1
done
