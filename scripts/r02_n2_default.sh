cd $GRAFT_REPO_ROOT
T=${1:-r02v}
timeout 900 python -m pytest tests -m gpu -q -x -k "colsum or reduce or unshuffle or engine or mae_base_all or classifier_against_oracle" 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/${T}_bench_n2.json 2> gpurun_out/${T}_bench_n2.err; echo "bench n2 rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_bench_n2.json').read().strip().splitlines()[-1])
print(round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'], d['dp_check'])
for k,v in d['extra_configs'].items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ('value','ms_per_step','step_frac_of_peak','error')})
PY
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 | tail -1 | cut -c1-300
