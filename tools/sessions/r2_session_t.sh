#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 120 python -m pytest tests/test_int8_gpu.py -m gpu -q -x -k "prefill_shapes" > $O/t_pytest_a.log 2>&1; echo "exit $?" >> $O/t_pytest_a.log; tail -4 $O/t_pytest_a.log
timeout 300 python -m pytest tests/test_int8_gpu.py tests/test_fp8_gpu.py tests/test_fullsize_gpu.py tests/test_quant_golden_gpu.py -m gpu -q -x > $O/t_pytest.log 2>&1; echo "exit $?" >> $O/t_pytest.log; tail -4 $O/t_pytest.log
for pair in 1 0; do
  for m in 2048 8192; do
    B200_GEMM_PAIR=$pair timeout 100 python tools/kbench.py gemm_int8 --m $m 2>&1 | sed "s/^{/{\"pair\": $pair, /" | cut -c1-200 | tee -a $O/t_kbench_int8_pair.jsonl
  done
done
B200_GEMM_PAIR=1 timeout 100 python tools/kbench.py gemm_fp8 --chan --m 8192 2>&1 | cut -c1-260 | tee $O/t_kbench_fp8_chan_pair.jsonl
