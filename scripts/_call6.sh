set -x
python scripts/knn_probe.py 50000x10000 3300x4300 > gpurun_out/plain6.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 12 -c 4 --csv --log-file gpurun_out/knn_launches3.csv python scripts/knn_probe.py 50000x10000 > gpurun_out/ncu6.log 2>&1
cat gpurun_out/plain6.log
python scripts/ransac_probe.py 100000 2 5 2>&1 | tee gpurun_out/r02_ransac_probe_st6.log
python -m pytest tests/test_gpu_ransac.py -q -x 2>&1 | tail -3
