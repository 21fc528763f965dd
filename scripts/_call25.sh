set -x
# final profiles of the shipped kernels
python scripts/knn_probe.py 50000x10000 > gpurun_out/plain25a.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"knn_tc_kernel|knn_rerank" -s 6 -c 2 -o gpurun_out/r02_knn_final python scripts/knn_probe.py 50000x10000 > gpurun_out/ncu25a.log 2>&1
python scripts/ransac_probe.py 100000 2 1 > gpurun_out/plain25b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:ransac_tc_score -s 5 -c 2 -o gpurun_out/r02_score_final python scripts/ransac_probe.py 100000 2 1 > gpurun_out/ncu25b.log 2>&1
CB_BENCH_WORKERS=1 CB_SYM_STREAMS=0 python bench.py --steps 2 --warmup 1 --quick > gpurun_out/plain25c.log 2>&1 &&
CB_BENCH_WORKERS=1 CB_SYM_STREAMS=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 1 --quick > gpurun_out/ncu25c.log 2>&1
tail -2 gpurun_out/ncu25a.log gpurun_out/ncu25b.log; wc -l gpurun_out/r02_launches.csv
