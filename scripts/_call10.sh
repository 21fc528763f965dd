python -m pytest tests/test_gpu_evaluation.py -x -q --tb=short 2>&1 | grep -v Warning | grep -v "^  " | head -40
