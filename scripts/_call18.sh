for v in 1 3 4; do echo "float share $v/4:"; CORSAIR_B200_LIB=$PWD/corsair_b200/lib/libcb_fs$v.so python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1; done
echo "float share 2/4:"; python scripts/ransac_probe.py 100000 2 5 2>&1 | tail -1
CORSAIR_B200_LIB=$PWD/corsair_b200/lib/libcb_fs3.so python -m pytest tests/test_gpu_ransac.py -x -q 2>&1 | tail -2
