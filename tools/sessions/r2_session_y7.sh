#!/bin/bash
# session Y7: block-FP8 stream-K with the contribution flag released by an idle warp
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_fp8_gpu.py tests/test_fullsize_gpu.py -m gpu -q -x > $O/y7_pytest_fp8.log 2>&1; echo "pytest exit $?" >> $O/y7_pytest_fp8.log
tail -3 $O/y7_pytest_fp8.log
rm -f $O/y7_kbench_fp8_block.jsonl
for m in 16 32 64 128; do
  timeout 300 python tools/kbench.py gemm_fp8 --m $m 2>/dev/null >> $O/y7_kbench_fp8_block.jsonl
done
cut -c1-130 $O/y7_kbench_fp8_block.jsonl
python tools/gemm_trace_cta.py 64 28672 4096 > $O/y7_trace_cta.txt 2>&1; tail -6 $O/y7_trace_cta.txt
timeout 600 python bench.py --quant fp8 --no-cpu-baseline > $O/y7_bench_decode_fp8.log 2> $O/y7_bench_decode_fp8.err; echo "exit $?" >> $O/y7_bench_decode_fp8.err
tail -n 1 $O/y7_bench_decode_fp8.log | cut -c1-330
