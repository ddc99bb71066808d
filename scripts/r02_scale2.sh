# usage: bash scripts/r02_scale2.sh TAG N : A/B of the stream schedule at N GPUs
cd $GRAFT_REPO_ROOT
T=${1:-r02d}; N=${2:-2}
run() { n=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 30 --warmup 5 --extra none --no-cpu-baseline > gpurun_out/${T}_$n.json 2> gpurun_out/${T}_$n.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${T}_$n.json').read().strip().splitlines()[-1])
    print('$n N=$N', round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'], d['dp_check'])
except Exception as e:
    print('$n', 'FAILED', e); print(open('gpurun_out/${T}_$n.err').read()[-1500:])
PY
}
run base SSL4P_WGRAD_OVERLAP=0 SSL4P_BUCKET_ADAMW=0
run both SSL4P_WGRAD_OVERLAP=1 SSL4P_BUCKET_ADAMW=1
run adamw SSL4P_WGRAD_OVERLAP=0 SSL4P_BUCKET_ADAMW=1
