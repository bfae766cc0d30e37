#!/bin/bash
# round-2 GPU session ZC: regression after the block-FP8 decode GEMM / TP exchange changes — GPU suite, smoke, headline,
# block-FP8 decode variant, the default-steps reference arm
set -u
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q > $O/zc_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/zc_pytest_gpu.log; tail -4 $O/zc_pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $O/zc_smoke.log 2>&1; echo "smoke exit $?" >> $O/zc_smoke.log; tail -2 $O/zc_smoke.log
timeout 600 python bench.py > $O/zc_bench.log 2> $O/zc_bench.err; echo "bench exit $?" >> $O/zc_bench.err; tail -n 1 $O/zc_bench.log | cut -c1-400
timeout 600 python bench.py --quant fp8 --no-cpu-baseline > $O/zc_bench_decode_fp8.log 2> $O/zc_bench_decode_fp8.err; tail -n 1 $O/zc_bench_decode_fp8.log | cut -c1-300
timeout 600 python bench.py --workload prefill --quant fp8 --no-cpu-baseline > $O/zc_bench_prefill_fp8.log 2> $O/zc_bench_prefill_fp8.err; tail -n 1 $O/zc_bench_prefill_fp8.log | cut -c1-300
