# usage: bash scripts/r04_n2.sh TAG : on 2 GPUs - the engine / capture tests, then the driver's launch line at N = 2 (NCCL inside the capture)
cd $GRAFT_REPO_ROOT
T=${1:-r04h}
timeout 600 python -m pytest tests -m gpu -q -x -k "graph or engine or pretrain_cli or classification_cli" 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/${T}_bench_n2.json 2> gpurun_out/${T}_bench_n2.err; echo "bench n2 rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${T}_bench_n2.json').read().strip().splitlines()[-1])
    print(round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'], d['dp_check'], 'frac', round(d['roofline']['frac'],3))
    for k,v in d['extra_configs'].items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ('value','ms_per_step','step_frac_of_peak','error')})
except Exception as e:
    print('FAILED', e); print(open('gpurun_out/${T}_bench_n2.err').read()[-3000:])
PY
