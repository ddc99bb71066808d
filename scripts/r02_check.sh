# usage: bash scripts/r02_check.sh TAG   (one B200): GPU tests, smoke, bench line
cd $GRAFT_REPO_ROOT
T=${1:-r02a}
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -25 gpurun_out/${T}_pytest.log
python __graft_entry__.py smoke 2>&1 | tail -1
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/${T}_bench.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_bench.json').read().strip().splitlines()[-1])
print(round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'], 'gemm frac', round(d['roofline']['frac'],3), 'step frac', round(d['roofline']['step_frac_of_peak'],3))
for k,v in d['extra_configs'].items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ('value','ms_per_step','step_frac_of_peak','error')})
PY
