set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python scripts/knn_probe.py 50000x10000 10000x10000 3300x4300 2>&1 | tee gpurun_out/r02_knn_probe_v3.log
python bench.py --steps 4 --warmup 2 --quick 2>&1 | tail -3 | tee gpurun_out/r02_bench_quick2.log
