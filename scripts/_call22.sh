python scripts/knn_probe.py 50000x10000 > gpurun_out/plain22.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 12 -c 8 --csv --log-file gpurun_out/knn_launches5.csv python scripts/knn_probe.py 50000x10000 > gpurun_out/ncu22.log 2>&1
cat gpurun_out/plain22.log
