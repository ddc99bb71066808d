# Round-end evidence run on one B200: parity tests, bench (both arms), ncu launch list, DRAM traffic of every GEMM launch,
# full-set captures of the dominant kernels.  Outputs under gpurun_out/ with the tag given as $1.
set -x
cd $GRAFT_REPO_ROOT
T=${1:-r01g}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -3 gpurun_out/${T}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e > gpurun_out/${T}_ncu.log 2>&1; echo "ncu rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:gemm_tcgen05 --launch-skip 496 -c 248 --csv --log-file gpurun_out/${T}_gemm_traffic.csv python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e > gpurun_out/${T}_ncu2.log 2>&1; echo "ncu traffic rc=$?"
GEMM_ONLY="dec fc1,dec dfc2,dec qkv,enc wfc1" timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -c 8 -o gpurun_out/${T}_gemm_full python scripts/gemm_bench.py 2 > gpurun_out/${T}_ncu3.log 2>&1; echo "ncu gemm full rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:attention --launch-skip 2 -c 2 -o gpurun_out/${T}_attn_full python scripts/attn_one.py > gpurun_out/${T}_ncu4.log 2>&1; echo "ncu attn full rc=$?"
ls -la gpurun_out | tail -15
