#!/bin/bash
# round-2 GPU session F: new two-tile TMA extend kernel + token-block stream-K GEMM — parity first, then timing
set -u
O=gpurun_out; mkdir -p $O
timeout 300 python -m pytest tests/test_attention_gpu.py -m gpu -q -x -k "extend or chunked" > $O/f_pytest_extend.log 2>&1; echo "exit $?" >> $O/f_pytest_extend.log; tail -15 $O/f_pytest_extend.log
timeout 300 python -m pytest tests/test_fullsize_gpu.py -m gpu -q -k extend > $O/f_pytest_fullsize.log 2>&1; echo "exit $?" >> $O/f_pytest_fullsize.log; tail -5 $O/f_pytest_fullsize.log
timeout 600 python -m pytest tests/test_int8_gpu.py -m gpu -q -x > $O/f_pytest_int8.log 2>&1; echo "exit $?" >> $O/f_pytest_int8.log; tail -8 $O/f_pytest_int8.log
for fa in 1 0; do
  B200_EXTEND_FA=$fa timeout 120 python tools/kbench.py extend --ctx 8192 2>&1 | tee -a $O/f_kbench_extend_fa$fa.jsonl
  B200_EXTEND_FA=$fa timeout 120 python tools/kbench.py extend --ctx 8192 --chunk 2048 2>&1 | tee -a $O/f_kbench_extend_fa$fa.jsonl
done
for tp in 4 8 2 1; do
  for bn in 64 128 256; do
    for m in 128 256 512; do
      B200_SK_BN=$bn timeout 120 python tools/kbench.py gemm_int8 --tp $tp --m $m 2>&1 | grep "layer total" | tee -a $O/f_kbench_int8_tokenblocks.jsonl
    done
  done
done
