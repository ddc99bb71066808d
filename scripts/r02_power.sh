cd $GRAFT_REPO_ROOT
T=${1:-r02c}
for steps in 20 200; do
for ov in 0 1; do
SSL4P_WGRAD_OVERLAP=$ov SSL4P_BUCKET_ADAMW=$ov timeout 300 python bench.py --steps $steps --warmup 5 --extra none --no-cpu-baseline --no-e2e > gpurun_out/${T}_s${steps}_o$ov.json 2> gpurun_out/${T}_s${steps}_o$ov.err
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_s${steps}_o$ov.json').read().strip().splitlines()[-1])
print('steps $steps overlap $ov', round(d['value']), round(d['ms_per_step'],2), d['clocks'])
PY
done; done
nvidia-smi -q -d POWER | grep -i -E "power limit|draw|cap" | head
