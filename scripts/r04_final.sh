# usage: bash scripts/r04_final.sh TAG : round-end evidence on one B200: every GPU test, smoke, bench (both arms), input / augment
# benches, ncu launch list of the step, DRAM traffic of every GEMM launch, ncu --set full summaries (reports summarised on the box)
cd $GRAFT_REPO_ROOT
T=${1:-r04f}
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -4 gpurun_out/${T}_pytest.log
python __graft_entry__.py smoke 2>&1 | tail -1
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_bench.json').read().strip().splitlines()[-1])
print(round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']), d['clocks']['sm_mhz'], 'gemm frac', round(d['roofline']['frac'],3), 'gemm ms', round(d['roofline']['gemm_ms_per_step'],2), 'step frac', round(d['roofline']['step_frac_of_peak'],3), 'launches', d['gpu_launches'])
for k,v in d['extra_configs'].items(): print(k, {a:(round(b,3) if isinstance(b,float) else b) for a,b in v.items() if a in ('value','ms_per_step','step_frac_of_peak','error')})
r=json.loads(open('gpurun_out/${T}_bench_ref.json').read().strip().splitlines()[-1]); print('reference arm', r.get('value'), r.get('unit'), r.get('cpu_baseline',{}).get('cores'))
PY
timeout 300 python scripts/augment_bench.py > gpurun_out/${T}_augment_bench_576x720.json 2>/dev/null; echo "augment rc=$?"
timeout 300 python scripts/input_bench.py 2>/dev/null | tail -1 > gpurun_out/${T}_input_bench_576x720.json; echo "input rc=$?"
timeout 300 python scripts/attn_bench.py 5 > gpurun_out/${T}_attn_bench.txt 2>&1; cat gpurun_out/${T}_attn_bench.txt
timeout 300 python scripts/gemm_bench.py 5 > gpurun_out/${T}_gemm_bench.txt 2>&1; tail -1 gpurun_out/${T}_gemm_bench.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e --extra none > gpurun_out/${T}_ncu.log 2>&1; echo "ncu launches rc=$?"
python scripts/ncu_summary.py gpurun_out/${T}_launches.csv > gpurun_out/${T}_launches_summary.txt
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:gemm_tcgen05 --launch-skip 496 -c 248 --csv --log-file gpurun_out/${T}_gemm_traffic.csv python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e --extra none > gpurun_out/${T}_ncu2.log 2>&1; echo "ncu traffic rc=$?"
python scripts/gemm_traffic_json.py gpurun_out/${T}_gemm_traffic.csv gpurun_out/${T}_gemm_traffic.json
GEMM_ONLY="dec fc1,dec dfc2,dec qkv,enc wfc1,dec dfc1" timeout 300 ncu --set full --clock-control none -k regex:gemm_tcgen05 -c 10 -o /tmp/${T}_gemm_full python scripts/gemm_bench.py 2 > gpurun_out/${T}_ncu3.log 2>&1; echo "ncu gemm full rc=$?"
python scripts/ncu_full_summary.py /tmp/${T}_gemm_full.ncu-rep > gpurun_out/${T}_gemm_full_summary.txt
du -sh gpurun_out
