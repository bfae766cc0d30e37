#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_int8_gpu.py tests/test_fp8_gpu.py -x -q > $O/d_pytest.log 2>&1; echo "pytest exit $?" >> $O/d_pytest.log
tail -3 $O/d_pytest.log
timeout 600 python tools/gemm_trace.py 64 > $O/d_trace.log 2>&1; grep streamk $O/d_trace.log | cut -c1-800
rm -f $O/d_kbench.log
for m in 64 128 256; do timeout 300 python tools/kbench.py gemm_int8 --m $m >> $O/d_kbench.log 2>&1; done
cut -c1-200 $O/d_kbench.log
timeout 900 python bench.py --no-cpu-baseline > $O/d_bench.log 2> $O/d_bench.err; tail -n 1 $O/d_bench.log | cut -c1-400; tail -3 $O/d_bench.err
