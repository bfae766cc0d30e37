#!/bin/bash
# session Y4: block-FP8 stream-K after the scale loads were made asynchronous; one vs two promotion warps per quarter
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_fp8_gpu.py tests/test_fullsize_gpu.py tests/test_elementwise_gpu.py -m gpu -q -x > $O/y4_pytest_fp8.log 2>&1; echo "pytest exit $?" >> $O/y4_pytest_fp8.log
tail -3 $O/y4_pytest_fp8.log
rm -f $O/y4_kbench_fp8_block_ew.jsonl
for ew in 4 8; do
  for m in 16 32 64 128; do
    B200_BLK_SK_EW=$ew timeout 300 python tools/kbench.py gemm_fp8 --m $m 2>/dev/null | sed "s/^{/{\"ew\": $ew, /" >> $O/y4_kbench_fp8_block_ew.jsonl
  done
done
cut -c1-140 $O/y4_kbench_fp8_block_ew.jsonl
python tools/gemm_trace.py 64 128 block > $O/y4_gemm_trace_block.jsonl 2>/dev/null
timeout 600 python bench.py --quant fp8 --no-cpu-baseline > $O/y4_bench_decode_fp8.log 2> $O/y4_bench_decode_fp8.err; echo "exit $?" >> $O/y4_bench_decode_fp8.err
tail -n 1 $O/y4_bench_decode_fp8.log | cut -c1-330
