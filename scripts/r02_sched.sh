# usage: bash scripts/r02_sched.sh TAG : dynamic GEMM tile scheduler on one B200: GEMM tests (timeout-guarded), per-shape table, bench A/B
cd $GRAFT_REPO_ROOT
T=${1:-r02i}
timeout 400 python -m pytest tests/test_ops_gpu.py -m gpu -q -x -k "gemm" > gpurun_out/${T}_gemm_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_gemm_pytest.log
tail -5 gpurun_out/${T}_gemm_pytest.log
SSL4P_GEMM_SCHED=0 timeout 300 python scripts/gemm_bench.py 5 2>&1 | tail -1
SSL4P_GEMM_SCHED=1 timeout 300 python scripts/gemm_bench.py 5 > gpurun_out/${T}_gemm_bench.txt 2>&1; tail -1 gpurun_out/${T}_gemm_bench.txt
run() { n=$1; shift
  env "$@" timeout 300 python bench.py --steps 30 --warmup 5 --extra none --no-cpu-baseline > gpurun_out/${T}_$n.json 2> gpurun_out/${T}_$n.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${T}_$n.json').read().strip().splitlines()[-1])
    print('$n', round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'], 'gemm ms', round(d['roofline']['gemm_ms_per_step'],2), 'loss', d['loss'])
except Exception as e:
    print('$n', 'FAILED', e); print(open('gpurun_out/${T}_$n.err').read()[-2500:])
PY
}
run sched0 SSL4P_GEMM_SCHED=0
run sched1 SSL4P_GEMM_SCHED=1
run sched0b SSL4P_GEMM_SCHED=0
run sched1b SSL4P_GEMM_SCHED=1
