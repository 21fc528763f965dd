CORSAIR_B200_LIB=$PWD/corsair_b200/lib/libcb_ph.so timeout 300 python -m pytest tests/test_gpu_ransac.py -x -q 2>&1 | tail -2
echo "phased (128 hyps/CTA, 4 TMEM stages):"; CORSAIR_B200_LIB=$PWD/corsair_b200/lib/libcb_ph.so timeout 120 python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
echo "default:"; python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
echo "phased:"; CORSAIR_B200_LIB=$PWD/corsair_b200/lib/libcb_ph.so timeout 120 python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
echo "default:"; python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
