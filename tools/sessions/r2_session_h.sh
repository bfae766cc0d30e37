#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 200 python -m pytest tests/test_attention_gpu.py -m gpu -q -x -k "extend or chunked" > $O/h_pytest_extend.log 2>&1; echo "exit $?" >> $O/h_pytest_extend.log; tail -6 $O/h_pytest_extend.log
timeout 200 python -m pytest tests/test_fullsize_gpu.py -m gpu -q -k extend > $O/h_pytest_fullsize.log 2>&1; echo "exit $?" >> $O/h_pytest_fullsize.log; tail -4 $O/h_pytest_fullsize.log
for poly in 2 0 4 6; do
  B200_FA_POLY=$poly timeout 60 python tools/kbench.py extend --ctx 8192 2>&1 | sed "s/^{/{\"poly\": $poly, /" | tee -a $O/h_kbench_extend_fa.jsonl
  B200_FA_POLY=$poly timeout 60 python tools/kbench.py extend --ctx 8192 --chunk 2048 2>&1 | sed "s/^{/{\"poly\": $poly, /" | tee -a $O/h_kbench_extend_fa.jsonl
done
