set -x
python scripts/knn_probe.py 50000x10000 10000x10000 3300x4300 2>&1 | tee gpurun_out/r02_knn_probe_v2.log
python scripts/ransac_probe.py 100000 2 5 2>&1 | tee gpurun_out/r02_ransac_probe_rot.log
python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_knn.py -q 2>&1 | tail -8
python scripts/knn_probe.py 50000x10000 > gpurun_out/plain4.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 12 -c 8 --csv --log-file gpurun_out/knn_launches2.csv python scripts/knn_probe.py 50000x10000 > gpurun_out/ncu4.log 2>&1
