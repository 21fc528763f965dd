python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_knn.py tests/test_gpu_golden.py -q 2>&1 | tail -3
python scripts/knn_probe.py 50000x10000 10000x10000 3300x4300 2>&1
python scripts/knn_probe.py 50000x10000 > gpurun_out/plain24.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 12 -c 4 --csv --log-file gpurun_out/knn_launches6.csv python scripts/knn_probe.py 50000x10000 > gpurun_out/ncu24.log 2>&1
