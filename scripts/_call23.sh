echo "two CTAs per SM (kHT=1):"; CORSAIR_B200_LIB=$PWD/corsair_b200/lib/libcb_ht1.so python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
echo "one CTA per SM (kHT=2):"; python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
CORSAIR_B200_LIB=$PWD/corsair_b200/lib/libcb_ht1.so python -m pytest tests/test_gpu_ransac.py -x -q 2>&1 | tail -2
echo "two CTAs per SM (kHT=1):"; CORSAIR_B200_LIB=$PWD/corsair_b200/lib/libcb_ht1.so python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
echo "one CTA per SM (kHT=2):"; python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
