#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_int8_gpu.py tests/test_fp8_gpu.py -x -q > $O/c_pytest.log 2>&1; echo "pytest exit $?" >> $O/c_pytest.log
tail -4 $O/c_pytest.log
timeout 600 python tools/gemm_trace.py 64 > $O/c_trace.log 2>&1; grep streamk $O/c_trace.log | cut -c1-700
rm -f $O/c_kbench.log
for m in 64 128 256; do timeout 300 python tools/kbench.py gemm_int8 --m $m >> $O/c_kbench.log 2>&1; done
cut -c1-200 $O/c_kbench.log
timeout 900 python bench.py --no-cpu-baseline > $O/c_bench.log 2> $O/c_bench.err; tail -n 1 $O/c_bench.log | cut -c1-400; tail -3 $O/c_bench.err
B200_GEMM_STREAMK=0 timeout 900 python bench.py --no-cpu-baseline > $O/c_bench_cluster.log 2> $O/c_bench_cluster.err; tail -n 1 $O/c_bench_cluster.log | cut -c1-400
