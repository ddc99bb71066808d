cd $GRAFT_REPO_ROOT
T=${1:-r02s}
timeout 400 python -m pytest tests/test_ops_gpu.py -m gpu -q -x -k "gemm or colsum" 2>&1 | tail -2
for pf in 0 1 0 1; do
echo "resid prefetch $pf"; SSL4P_GEMM_RESID_PREFETCH=$pf GEMM_ONLY="enc proj,enc fc2,dec proj,dec fc2" timeout 300 python scripts/gemm_bench.py 5 2>&1 | grep -E "proj|fc2"
done
runb() { n=$1; shift
  env "$@" timeout 300 python bench.py --steps 30 --warmup 5 --extra none --no-cpu-baseline --no-e2e > gpurun_out/${T}_mae_$n.json 2> gpurun_out/${T}_mae_$n.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/${T}_mae_$n.json').read().strip().splitlines()[-1])
print('mae $n', round(d['value']), round(d['ms_per_step'],3), d['clocks']['sm_mhz'])
PY
}
runb pf0_a SSL4P_GEMM_RESID_PREFETCH=0
runb pf1_a SSL4P_GEMM_RESID_PREFETCH=1
runb pf0_b SSL4P_GEMM_RESID_PREFETCH=0
runb pf1_b SSL4P_GEMM_RESID_PREFETCH=1
