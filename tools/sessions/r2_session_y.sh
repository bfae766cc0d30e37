#!/bin/bash
# round-2 GPU session Y: block-FP8 stream-K decode GEMM, one vs two promotion warps per TMEM lane quarter
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_fp8_gpu.py tests/test_fullsize_gpu.py -m gpu -q -x > $O/y_pytest_fp8.log 2>&1; echo "pytest exit $?" >> $O/y_pytest_fp8.log
tail -4 $O/y_pytest_fp8.log
for ew in 4 8; do
  for m in 16 32 64 128; do
    B200_BLK_SK_EW=$ew timeout 300 python tools/kbench.py gemm_fp8 --m $m 2>/dev/null | sed "s/^{/{\"ew\": $ew, /" >> $O/y_kbench_fp8_block_ew.jsonl
  done
done
cut -c1-160 $O/y_kbench_fp8_block_ew.jsonl
timeout 600 python bench.py --quant fp8 --no-cpu-baseline > $O/y_bench_decode_fp8.log 2> $O/y_bench_decode_fp8.err; echo "exit $?" >> $O/y_bench_decode_fp8.err
tail -n 1 $O/y_bench_decode_fp8.log | cut -c1-400
