#!/bin/bash
# Session H: FP8-KV tests first (new code), its kernel bench, then the whole GPU suite, bench and launch list.
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_fp8kv_gpu.py -x -q > $O/h_fp8kv.log 2>&1; echo "exit $?" >> $O/h_fp8kv.log; tail -15 $O/h_fp8kv.log
timeout 300 python tools/kbench.py decode --fp8kv > $O/h_kbench_fp8kv.log 2>&1; tail -6 $O/h_kbench_fp8kv.log
timeout 300 python tools/kbench.py decode --chunk 0 > $O/h_kbench_bf16.log 2>&1; tail -2 $O/h_kbench_bf16.log
timeout 2400 python -m pytest tests -m gpu -x -q > $O/h_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/h_pytest_gpu.log; tail -5 $O/h_pytest_gpu.log
timeout 900 python bench.py > $O/h_bench.log 2> $O/h_bench.err; echo "bench exit $?" >> $O/h_bench.err; tail -n 1 $O/h_bench.log | cut -c1-2500; tail -n 2 $O/h_bench.err
