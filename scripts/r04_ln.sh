# usage: bash scripts/r04_ln.sh TAG : no-alloc entry point test (alone), LayerNorm tests under both backward variants, then
# step A/B of the shared-memory LayerNorm backward (SSL4P_LN_BWD_SMEM=0/1) on the classifier and MAE steps, and an ncu timing
cd $GRAFT_REPO_ROOT
T=${1:-r04a}
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -15
SSL4P_LN_BWD_SMEM=1 timeout 600 python -m pytest tests -m gpu -q -x -k "layernorm or mae_base or mae_large or classifier" 2>&1 | tail -4
SSL4P_LN_BWD_SMEM=0 timeout 300 python -m pytest tests -m gpu -q -x -k "layernorm" 2>&1 | tail -2
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
runb() { n=$1; shift
  env "$@" timeout 300 python bench.py --steps 30 --warmup 5 --extra none --no-cpu-baseline --no-e2e > gpurun_out/${T}_mae_$n.json 2> gpurun_out/${T}_mae_$n.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${T}_mae_$n.json').read().strip().splitlines()[-1])
    print('mae $n', round(d['value']), round(d['ms_per_step'],3), d['clocks']['sm_mhz'])
except Exception as e:
    print('$n', 'FAILED', e); print(open('gpurun_out/${T}_mae_$n.err').read()[-1500:])
PY
}
run reg SSL4P_LN_BWD_SMEM=0
run smem SSL4P_LN_BWD_SMEM=1
run reg2 SSL4P_LN_BWD_SMEM=0
run smem2 SSL4P_LN_BWD_SMEM=1
runb reg SSL4P_LN_BWD_SMEM=0
runb smem SSL4P_LN_BWD_SMEM=1
runb reg2 SSL4P_LN_BWD_SMEM=0
runb smem2 SSL4P_LN_BWD_SMEM=1
for v in 0 1; do
SSL4P_LN_BWD_SMEM=$v timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:layernorm_bwd -c 60 --csv --log-file gpurun_out/${T}_ln_$v.csv python scripts/bench_classifier.py --steps 1 --warmup 1 > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open('gpurun_out/${T}_ln_$v.csv')) if len(r)>10 and r[0].isdigit()]
t={}
for r in rows:
    t.setdefault((r[4][:40], r[-3]), []).append(float(r[-1].replace(',','')))
for k,v in t.items(): print('ln smem=$v', k, len(v), sum(v)/len(v))
PY
done
