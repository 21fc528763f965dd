python -m pytest tests/test_gpu_ransac.py tests/test_gpu_knn_tc.py -x -q 2>&1 | tail -3
CB_RT_EPI_WARPS=8 python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
python scripts/knn_probe.py 50000x10000 10000x10000 3300x4300 2>&1
