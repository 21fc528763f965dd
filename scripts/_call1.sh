set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
python scripts/knn_probe.py 50000x10000 10000x10000 3300x4300 > gpurun_out/r02_knn_probe_base.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"knn_tc_kernel|knn_rerank" -s 6 -c 2 -o gpurun_out/r02_knn_base python scripts/knn_probe.py 50000x10000 > gpurun_out/ncu_knn.log 2>&1
python scripts/ransac_probe.py 100000 2 5 > gpurun_out/r02_ransac_probe_base.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:ransac_tc_score -s 5 -c 1 -o gpurun_out/r02_score_base python scripts/ransac_probe.py 100000 2 1 > gpurun_out/ncu_score.log 2>&1
cat gpurun_out/r02_knn_probe_base.log gpurun_out/r02_ransac_probe_base.log
tail -3 gpurun_out/ncu_knn.log gpurun_out/ncu_score.log
