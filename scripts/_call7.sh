set -x
python scripts/knn_probe.py 50000x10000 > gpurun_out/plain7.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:knn_tc_kernel -s 6 -c 1 -o gpurun_out/r02_knn_v2 python scripts/knn_probe.py 50000x10000 > gpurun_out/ncu7.log 2>&1
tail -3 gpurun_out/ncu7.log
