python -m pytest tests/test_gpu_knn.py tests/test_gpu_knn_tc.py tests/test_gpu_golden.py -q 2>&1 | tail -3
python scripts/knn_probe.py 50000x10000 3300x4300
KNN_D=16 python scripts/knn_probe.py 50000x10000 3300x4300
python scripts/sanitize_cases.py 2>&1 | tail -7
