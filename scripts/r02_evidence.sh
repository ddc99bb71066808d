# usage: bash scripts/r02_evidence.sh TAG : one B200: GPU tests, smoke, bench, ncu launch list of a step, ncu --set full of
# the HBM-bound kernels (LayerNorm, loss, AdamW, token kernels) and of the epilogue-bound decoder GEMMs
cd $GRAFT_REPO_ROOT
T=${1:-r02c}
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -4 gpurun_out/${T}_pytest.log
python __graft_entry__.py smoke 2>&1 | tail -1
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_bench.json').read().strip().splitlines()[-1])
print(round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'], 'gemm frac', round(d['roofline']['frac'],3), 'gemm ms', round(d['roofline']['gemm_ms_per_step'],2), 'step frac', round(d['roofline']['step_frac_of_peak'],3))
for k,v in d['extra_configs'].items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ('value','ms_per_step','step_frac_of_peak','error')})
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e --extra none > gpurun_out/${T}_ncu.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"layernorm|masked_mse_fwd|masked_mse_bwd|adamw_flat|decoder_unshuffle|im2col|colsum_partial|splitk_reduce" --launch-skip 700 -c 60 -o gpurun_out/${T}_small_full python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e --extra none > gpurun_out/${T}_ncu2.log 2>&1; echo "ncu small full rc=$?"
GEMM_ONLY="dec fc1,dec dfc2,dec proj,dec dfc1" timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -c 8 -o gpurun_out/${T}_gemm_full python scripts/gemm_bench.py 2 > gpurun_out/${T}_ncu3.log 2>&1; echo "ncu gemm full rc=$?"
ls -la gpurun_out | tail -12
