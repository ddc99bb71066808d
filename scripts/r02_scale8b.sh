# usage: bash scripts/r02_scale8b.sh TAG N : dynamic tile scheduler on / off at N GPUs, then N = 1 on the same box
cd $GRAFT_REPO_ROOT
T=${1:-r02q}; N=${2:-8}
run() { n=$1; shift
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 30 --warmup 5 --extra none --no-cpu-baseline > gpurun_out/${T}_$n.json 2> gpurun_out/${T}_$n.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${T}_$n.json').read().strip().splitlines()[-1])
    c=d['dp_check'] or {}
    print('$n N=$N', round(d['value']), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'], 'bit_equal', c.get('params_bit_equal_across_ranks'), 'cos', c.get('grad_min_cosine_dp_vs_single_gpu'))
except Exception as e:
    print('$n', 'FAILED', e); print(open('gpurun_out/${T}_$n.err', errors='replace').read()[-1500:])
PY
}
run sched0 SSL4P_GEMM_SCHED=0
run sched1 SSL4P_GEMM_SCHED=1
run sched0b SSL4P_GEMM_SCHED=0
run sched1b SSL4P_GEMM_SCHED=1
for s in 0 1; do
SSL4P_GEMM_SCHED=$s timeout 300 python bench.py --steps 30 --warmup 5 --extra none --no-cpu-baseline > gpurun_out/${T}_n1_s$s.json 2> gpurun_out/${T}_n1_s$s.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_n1_s$s.json').read().strip().splitlines()[-1])
print('n1 sched$s', round(d['value']), round(d['ms_per_step'],3), d['clocks']['sm_mhz'])
PY
done
