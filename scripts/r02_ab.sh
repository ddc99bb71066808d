# usage: bash scripts/r02_ab.sh TAG : A/B of the stream schedule on one B200 (bench headline only), then the GPU tests
cd $GRAFT_REPO_ROOT
T=${1:-r02b}
run() { # name, env...
  n=$1; shift
  env "$@" timeout 300 python bench.py --steps 20 --warmup 5 --extra none --no-cpu-baseline > gpurun_out/${T}_$n.json 2> gpurun_out/${T}_$n.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${T}_$n.json').read().strip().splitlines()[-1])
    print('$n', round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'], 'gemm frac', round(d['roofline']['frac'],3), 'loss', d['loss'])
except Exception as e:
    print('$n', 'FAILED', e); print(open('gpurun_out/${T}_$n.err').read()[-1500:])
PY
}
run base SSL4P_WGRAD_OVERLAP=0 SSL4P_BUCKET_ADAMW=0
run overlap SSL4P_WGRAD_OVERLAP=1 SSL4P_BUCKET_ADAMW=0
run both SSL4P_WGRAD_OVERLAP=1 SSL4P_BUCKET_ADAMW=1
run adamw SSL4P_WGRAD_OVERLAP=0 SSL4P_BUCKET_ADAMW=1
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -12 gpurun_out/${T}_pytest.log
