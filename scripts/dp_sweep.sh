# data-parallel overlap sweep: SMs reserved for NCCL (0 = shared SMs), bucket size, and the step without the collective
cd $GRAFT_REPO_ROOT
N=${1:-2}
i=0
for cfg in "0 64 0" "0 64 1" "0 16 0" "0 512 0" "2 64 0" "16 64 0"; do
  set -- $cfg; i=$((i+1))
  SSL4P_DP_SMS=$1 SSL4P_DP_BUCKET_MB=$2 SSL4P_DP_SKIP_COLLECTIVE=$3 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29540+i)) bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --no-e2e 2> gpurun_out/dp_sweep_$i.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('dp_sms=$1 bucket_mb=$2 skip=$3', d['n_gpus'], round(d['value']), d['ms_per_step'])"
done
