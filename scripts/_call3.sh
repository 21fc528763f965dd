set -x
python scripts/knn_probe.py 50000x10000 3300x4300 > gpurun_out/plain3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 16 --csv --log-file gpurun_out/knn_launches.csv python scripts/knn_probe.py 50000x10000 3300x4300 > gpurun_out/ncu3.log 2>&1
grep -E "knn|Kernel Name" gpurun_out/knn_launches.csv | cut -d, -f5,15 | head -40
