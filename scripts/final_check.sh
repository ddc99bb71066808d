# Last confirmation of a round on one B200: every GPU test, smoke, input-transform bench, bench line.  usage: bash scripts/final_check.sh TAG
cd $GRAFT_REPO_ROOT
T=${1:-r01m}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; tail -3 gpurun_out/${T}_pytest.log
python __graft_entry__.py smoke 2>&1 | tail -1
for f in 576x720 1072x1920; do
  timeout 300 python scripts/input_bench.py --frame $f 2>&1 | tail -1 > gpurun_out/${T}_input_bench_$f.json
  python -c "
import json; d=json.loads(open('gpurun_out/${T}_input_bench_$f.json').read()); print(d['frame'], round(d['device_ms_per_batch'],3), round(d['e2e_images_per_s']), round(d['cpu_baseline']['images_per_s']))"
done
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench.json 2>/dev/null
python -c "
import json; d=json.loads(open('gpurun_out/${T}_bench.json').read().strip().splitlines()[-1]); print(round(d['value']), round(d['ms_per_step'],2), round(d['e2e']['value']), d['clocks']['sm_mhz'], round(d['roofline']['frac'],3))"
