# usage: bash scripts/r04_extra_launches.sh TAG : ncu launch lists of the ViT-L/16 pretraining step and of the classifier fine-tune step on the final code
cd $GRAFT_REPO_ROOT
T=${1:-r04m}
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 1100 -c 1100 --csv --log-file gpurun_out/${T}_large_launches.csv python bench.py --model large --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e --extra none > gpurun_out/${T}_ncu_l.log 2>&1; echo "ncu large rc=$?"
python scripts/ncu_summary.py gpurun_out/${T}_large_launches.csv > gpurun_out/${T}_large_launches_summary.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 900 -c 900 --csv --log-file gpurun_out/${T}_cls_launches.csv python scripts/bench_classifier.py --steps 2 --warmup 2 > gpurun_out/${T}_ncu_c.log 2>&1; echo "ncu cls rc=$?"
python scripts/ncu_summary.py gpurun_out/${T}_cls_launches.csv > gpurun_out/${T}_cls_launches_summary.txt
head -12 gpurun_out/${T}_large_launches_summary.txt | cut -c1-110; head -14 gpurun_out/${T}_cls_launches_summary.txt | cut -c1-110
rm -f gpurun_out/${T}_large_launches.csv gpurun_out/${T}_cls_launches.csv
