# usage: bash scripts/r04_n4.sh TAG : the driver's launch line at N = 4 on the final code (NCCL all-reduces inside the thread_local capture)
cd $GRAFT_REPO_ROOT
T=${1:-r04k}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 4 --steps 20 --warmup 5 --extra none --no-cpu-baseline > gpurun_out/${T}_bench_n4.json 2> gpurun_out/${T}_bench_n4.err; echo "bench n4 rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${T}_bench_n4.json').read().strip().splitlines()[-1])
    print(round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'], d['dp_check'])
except Exception as e:
    print('FAILED', e); print(open('gpurun_out/${T}_bench_n4.err').read()[-3000:])
PY
