# usage: bash scripts/r03_symm.sh TAG N : gradient exchange through torch's symmetric-memory multimem all-reduce vs NCCL at N GPUs
cd $GRAFT_REPO_ROOT
T=${1:-r03j}; N=${2:-2}
run() { n=$1; shift
  env "$@" timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 30 --warmup 5 --extra none --no-cpu-baseline > gpurun_out/${T}_$n.json 2> gpurun_out/${T}_$n.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${T}_$n.json').read().strip().splitlines()[-1])
    c=d['dp_check'] or {}
    print('$n N=$N', round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'], 'bit_equal', c.get('params_bit_equal_across_ranks'), 'cos', c.get('grad_min_cosine_dp_vs_single_gpu'))
except Exception as e:
    print('$n', 'FAILED', e); print(open('gpurun_out/${T}_$n.err', errors='replace').read()[-2500:])
PY
}
run nccl SSL4P_DP_SYMM=0
run symm SSL4P_DP_SYMM=1
run nccl_b SSL4P_DP_SYMM=0
run symm_b SSL4P_DP_SYMM=1
timeout 300 python bench.py --steps 30 --warmup 5 --extra none --no-cpu-baseline > gpurun_out/${T}_n1.json 2> gpurun_out/${T}_n1.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_n1.json').read().strip().splitlines()[-1])
print('n1', round(d['value']), round(d['ms_per_step'],3), d['clocks']['sm_mhz'])
PY
