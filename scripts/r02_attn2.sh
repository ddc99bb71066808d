cd $GRAFT_REPO_ROOT
T=${1:-r02t}
timeout 600 python -m pytest tests/test_ops_gpu.py -m gpu -q -x -k "attention" > gpurun_out/${T}_attn_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_attn_pytest.log; tail -4 gpurun_out/${T}_attn_pytest.log
timeout 300 python scripts/attn_bench.py 5 2>&1 | tail -4
