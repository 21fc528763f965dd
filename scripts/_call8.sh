set -x
python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_knn.py -q 2>&1 | tail -8
python scripts/knn_probe.py 50000x10000 10000x10000 3300x4300 > gpurun_out/plain8.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 12 -c 4 --csv --log-file gpurun_out/knn_launches4.csv python scripts/knn_probe.py 50000x10000 > gpurun_out/ncu8.log 2>&1
cat gpurun_out/plain8.log
