set -x
python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_knn.py tests/test_gpu_golden.py -x -q 2>&1 | tail -15
python scripts/knn_probe.py 50000x10000 10000x10000 3300x4300 128x256 2>&1 | tee gpurun_out/r02_knn_probe_v1.log
