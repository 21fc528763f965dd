python -m pytest tests/test_gpu_shapenet.py tests/test_gpu_evaluation.py -q --tb=short 2>&1 | grep -v "Warning\|^  \|^$" | tail -60
