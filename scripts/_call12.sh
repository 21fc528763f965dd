CB_RT_EPI_WARPS=8 python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
python scripts/ransac_probe.py 100000 2 1 > gpurun_out/plain12.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:ransac_tc_score -s 5 -c 1 -o gpurun_out/r02_score_v2 python scripts/ransac_probe.py 100000 2 1 > gpurun_out/ncu12.log 2>&1
