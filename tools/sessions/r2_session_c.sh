#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_fp8_gpu.py tests/test_quant_golden_gpu.py -m gpu -q -x > $O/c_pytest_fp8.log 2>&1; echo "pytest exit $?" >> $O/c_pytest_fp8.log
tail -5 $O/c_pytest_fp8.log
for m in 64 128 8192; do
  for dbg in 0 1 3; do
    echo "== m=$m debug=$dbg"
    B200_GEMM_DEBUG=$dbg timeout 300 python tools/kbench.py gemm_fp8 --m $m 2>&1 | cut -c1-200 | tee -a $O/c_kbench_dbg.log
  done
done
