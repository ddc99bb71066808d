set -x
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r01c_pytest.log
tail -3 gpurun_out/r01c_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r01c_bench.json 2> gpurun_out/r01c_bench.err; echo "bench rc=$?"
cat gpurun_out/r01c_bench.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01c_bench_ref.json 2> gpurun_out/r01c_bench_ref.err; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r01c_launches.csv python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e > gpurun_out/r01c_ncu.log 2>&1; echo "ncu rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 --launch-skip 300 -c 12 -o gpurun_out/r01c_gemm_full python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e > gpurun_out/r01c_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out
