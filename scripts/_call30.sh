python -m pytest tests/test_gpu_knn_tc.py -q -k fallback_rows_split 2>&1 | tail -5
