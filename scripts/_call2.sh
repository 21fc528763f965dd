set -x
python -m pytest tests/test_gpu_knn_tc.py tests/test_gpu_knn.py tests/test_gpu_golden.py -x -q 2>&1 | tail -15
python scripts/knn_probe.py 50000x10000 10000x10000 3300x4300 128x256 2>&1 | tee gpurun_out/r02_knn_probe_v1.log
python -m pytest tests/test_gpu_ransac.py tests/test_gpu_symmetry.py -x -q 2>&1 | tail -15
CB_RT_EPI_WARPS=8 python scripts/ransac_probe.py 100000 2 5 2>&1 | tee gpurun_out/r02_ransac_probe_w8.log
python scripts/ransac_probe.py 100000 2 5 2>&1 | tee gpurun_out/r02_ransac_probe_w16.log
python bench.py --steps 3 --warmup 1 --quick 2>&1 | tail -5 | tee gpurun_out/r02_bench_quick1.log
