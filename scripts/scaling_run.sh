# bench.py at N GPUs of one box (the driver's launch line); usage: bash scripts/scaling_run.sh N TAG
cd $GRAFT_REPO_ROOT
N=$1; T=${2:-r01k}
if [ "$N" = "1" ]; then python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/${T}_scale_n1.json 2> gpurun_out/${T}_scale_n1.err
else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${T}_scale_n$N.json 2> gpurun_out/${T}_scale_n$N.err; fi
python -c "
import json,sys
d=json.loads(open('gpurun_out/${T}_scale_n$N.json').read().strip().splitlines()[-1])
print('N=%d  %.0f img/s  %.2f ms/step  e2e %s  clocks %s' % (d['n_gpus'], d['value'], d['ms_per_step'], d['e2e'] and round(d['e2e']['value']), d['clocks']['sm_mhz']))"
