python -m pytest tests/test_gpu_ransac.py -x -q 2>&1 | tail -2
CB_RT_EPI_WARPS=8 python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
