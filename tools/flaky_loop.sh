#!/bin/bash
# development helper: repeat the GPU parity file until the pipelined host-gemm test fails
for i in $(seq 1 ${1:-10}); do
  timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "${2:-gemm}" -p no:cacheprovider > gpurun_out/flaky_$i.log 2>&1
  rc=$?
  echo "run $i rc=$rc $(tail -1 gpurun_out/flaky_$i.log)"
  if [ $rc -ne 0 ]; then grep -n "assert\|Error\|bad" gpurun_out/flaky_$i.log | head -20; fi
done
