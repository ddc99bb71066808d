# usage: bash scripts/r04_ncu.sh TAG : one B200: ncu --set full of the HBM-bound kernels (LayerNorm incl. the shared-memory backward,
# loss, AdamW, token kernels, reductions) inside a real step, and of the attention kernels, with their hot SASS lines; summarised on the box
cd $GRAFT_REPO_ROOT
T=${1:-r04j}
timeout 900 ncu --set full --clock-control none -k regex:"layernorm|masked_mse_fwd|masked_mse_bwd|adamw_flat|decoder_unshuffle|im2col|colsum_partial|splitk_reduce|assemble_tokens|mask_token_grad|reduce_rows" --launch-skip 700 -c 60 -o /tmp/${T}_small_full python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e --extra none > gpurun_out/${T}_ncu2.log 2>&1; echo "ncu small full rc=$?"
python scripts/ncu_full_summary.py /tmp/${T}_small_full.ncu-rep > gpurun_out/${T}_small_full_summary.txt
timeout 300 ncu --set full --clock-control none --import-source on -k regex:attention --launch-skip 2 -c 2 -o /tmp/${T}_attn_full python scripts/attn_one.py > gpurun_out/${T}_ncu4.log 2>&1; echo "ncu attn full rc=$?"
python scripts/ncu_full_summary.py /tmp/${T}_attn_full.ncu-rep > gpurun_out/${T}_attn_full_summary.txt
for id in 0 1; do
  ncu -i /tmp/${T}_attn_full.ncu-rep --page source --csv --print-source sass --launch-skip $id --launch-count 1 > /tmp/asrc_$id.csv 2>/dev/null
  python scripts/ncu_mix.py /tmp/asrc_$id.csv 40 > gpurun_out/${T}_attn_hot_$id.txt 2>&1
done
du -sh gpurun_out
