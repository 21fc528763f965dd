#!/bin/bash
# Memory-safety evidence for the multi-role kernels on the small-shape cases of
# scripts/sanitize_cases.py (each case is also checked against the oracle):
#   1. plain run with the shipped library
#   2. the same run with the DEBUG library (make debug: device-side range checks that trap, common.cuh)
#   3. compute-sanitizer --tool <memcheck|racecheck|synccheck|initcheck>, ONE tool per gpurun call
#      (B200_PROFILING.md).  On the pool this repo was developed on the tool is closed by the operators
#      ("runs under it have left GPUs needing a reset"); the log then says so and steps 1-2 are the record.
tool=${1:-memcheck}
# the debug library is built here, not by __graft_entry__.build() (it doubles the nvcc time of every build)
make -C corsair_b200/csrc -j8 -s debug || exit 1
python scripts/sanitize_cases.py > gpurun_out/sanitize_plain.log 2>&1 || { tail -20 gpurun_out/sanitize_plain.log; exit 1; }
CORSAIR_B200_LIB=$PWD/corsair_b200/lib/libcorsair_b200_dbg.so python scripts/sanitize_cases.py > gpurun_out/sanitize_debuglib.log 2>&1
echo "debug library exit code $?" >> gpurun_out/sanitize_debuglib.log
timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python scripts/sanitize_cases.py > gpurun_out/sanitize_$tool.log 2>&1
echo "exit code $?" >> gpurun_out/sanitize_$tool.log
{ echo "== plain"; cat gpurun_out/sanitize_plain.log; echo "== debug library (CB_DEBUG_CHECKS)"; cat gpurun_out/sanitize_debuglib.log;
  echo "== compute-sanitizer --tool $tool"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|closed|exit code|Error" gpurun_out/sanitize_$tool.log | sort | uniq -c | sort -rn | head -30; } > gpurun_out/sanitize_$tool.txt
cat gpurun_out/sanitize_$tool.txt
