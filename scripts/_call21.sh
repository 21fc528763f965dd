python -m pytest tests/test_gpu_symmetry.py tests/test_gpu_golden.py tests/test_gpu_dist.py -q --tb=short -rs 2>&1 | grep -v "Warning\|^  \|^$" | tail -25
