# usage: bash scripts/r02_ncu.sh TAG : one B200: ncu launch list of a step, ncu --set full of the HBM-bound kernels and of the
# epilogue-bound decoder GEMMs; reports are summarised ON THE BOX (gpurun_out/ is capped at 64 MiB) and only the text comes back
cd $GRAFT_REPO_ROOT
T=${1:-r02d}
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e --extra none > gpurun_out/${T}_ncu.log 2>&1; echo "ncu launches rc=$?"
python scripts/ncu_summary.py gpurun_out/${T}_launches.csv > gpurun_out/${T}_launches_summary.txt
timeout 900 ncu --set full --clock-control none -k regex:"layernorm|masked_mse_fwd|masked_mse_bwd|adamw_flat|decoder_unshuffle|im2col|colsum_partial|splitk_reduce|assemble_tokens" --launch-skip 700 -c 40 -o /tmp/${T}_small_full python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e --extra none > gpurun_out/${T}_ncu2.log 2>&1; echo "ncu small full rc=$?"
python scripts/ncu_full_summary.py /tmp/${T}_small_full.ncu-rep > gpurun_out/${T}_small_full_summary.txt
GEMM_ONLY="dec fc1,dec dfc2,dec proj,dec dfc1" timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -c 8 -o /tmp/${T}_gemm_full python scripts/gemm_bench.py 2 > gpurun_out/${T}_ncu3.log 2>&1; echo "ncu gemm full rc=$?"
python scripts/ncu_full_summary.py /tmp/${T}_gemm_full.ncu-rep > gpurun_out/${T}_gemm_full_summary.txt
for id in 0 2 4 6; do
  ncu -i /tmp/${T}_gemm_full.ncu-rep --page source --csv --print-source sass --launch-skip $id --launch-count 1 > /tmp/src_$id.csv 2>/dev/null
  python scripts/ncu_mix.py /tmp/src_$id.csv 30 > gpurun_out/${T}_gemm_hot_$id.txt 2>&1
done
timeout 300 ncu --set full --clock-control none --import-source on -k regex:attention --launch-skip 2 -c 2 -o /tmp/${T}_attn_full python scripts/attn_one.py > gpurun_out/${T}_ncu4.log 2>&1; echo "ncu attn full rc=$?"
python scripts/ncu_full_summary.py /tmp/${T}_attn_full.ncu-rep > gpurun_out/${T}_attn_full_summary.txt
for id in 0 1; do
  ncu -i /tmp/${T}_attn_full.ncu-rep --page source --csv --print-source sass --launch-skip $id --launch-count 1 > /tmp/asrc_$id.csv 2>/dev/null
  python scripts/ncu_mix.py /tmp/asrc_$id.csv 40 > gpurun_out/${T}_attn_hot_$id.txt 2>&1
done
du -sh gpurun_out; ls -la gpurun_out | tail -20
