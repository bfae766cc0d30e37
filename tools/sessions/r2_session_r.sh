#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 300 python -m pytest tests/test_int8_gpu.py tests/test_fp8_gpu.py tests/test_fullsize_gpu.py -m gpu -q -x -k "prefill or m8192 or chan or fullsize or scaled_mm" > $O/r_pytest.log 2>&1; echo "exit $?" >> $O/r_pytest.log; tail -5 $O/r_pytest.log
for mc in 1 0; do
  for m in 2048 8192; do
    B200_GEMM_MC=$mc timeout 100 python tools/kbench.py gemm_int8 --m $m 2>&1 | sed "s/^{/{\"mc\": $mc, /" | cut -c1-200 | tee -a $O/r_kbench_int8_mc.jsonl
  done
done
timeout 100 python tools/kbench.py peak 2>&1 | cut -c1-300 | tee $O/r_kbench_peak.jsonl
