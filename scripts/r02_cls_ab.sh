# usage: bash scripts/r02_cls_ab.sh TAG : classifier fine-tune step (B=128) and ViT-L step under the GEMM switches, same box
cd $GRAFT_REPO_ROOT
T=${1:-r02l}
run() { n=$1; shift
  env "$@" timeout 300 python scripts/bench_classifier.py --steps 30 --warmup 5 > gpurun_out/${T}_cls_$n.json 2> gpurun_out/${T}_cls_$n.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${T}_cls_$n.json').read().strip().splitlines()[-1])
    print('cls $n', round(d['value']), round(d['ms_per_step'],3))
except Exception as e:
    print('$n', 'FAILED', e); print(open('gpurun_out/${T}_cls_$n.err').read()[-1500:])
PY
}
run sched0_cg0 SSL4P_GEMM_SCHED=0 SSL4P_GEMM_CG=0
run sched1_cg0 SSL4P_GEMM_SCHED=1 SSL4P_GEMM_CG=0
run sched0_cg1 SSL4P_GEMM_SCHED=0 SSL4P_GEMM_CG=1
run sched1_cg1 SSL4P_GEMM_SCHED=1 SSL4P_GEMM_CG=1
run sched1_cg2 SSL4P_GEMM_SCHED=1 SSL4P_GEMM_CG=2
run sched0_cg0b SSL4P_GEMM_SCHED=0 SSL4P_GEMM_CG=0
run sched1_cg0b SSL4P_GEMM_SCHED=1 SSL4P_GEMM_CG=0
runb() { n=$1; shift
  env "$@" timeout 300 python bench.py --steps 30 --warmup 5 --extra none --no-cpu-baseline --no-e2e > gpurun_out/${T}_mae_$n.json 2> gpurun_out/${T}_mae_$n.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_mae_$n.json').read().strip().splitlines()[-1])
print('mae $n', round(d['value']), round(d['ms_per_step'],3), d['clocks']['sm_mhz'])
PY
}
runb sched1_cg0 SSL4P_GEMM_SCHED=1 SSL4P_GEMM_CG=0
runb sched1_cg1 SSL4P_GEMM_SCHED=1 SSL4P_GEMM_CG=1
runb sched1_cg2 SSL4P_GEMM_SCHED=1 SSL4P_GEMM_CG=2
