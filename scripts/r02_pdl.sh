# usage: bash scripts/r02_pdl.sh TAG : A/B of programmatic dependent launch on one B200 (bench headline), then the GPU tests with it on
cd $GRAFT_REPO_ROOT
T=${1:-r02g}
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
run pdl0 SSL4P_PDL=0
run pdl1 SSL4P_PDL=1
run pdl0b SSL4P_PDL=0
run pdl1b SSL4P_PDL=1
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -6 gpurun_out/${T}_pytest.log
