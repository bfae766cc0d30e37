#!/bin/bash
# GPU session A: stream-K decode GEMM correctness + A/B against the cluster split-K kernel.
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_int8_gpu.py tests/test_fp8_gpu.py -x -q > $O/a_pytest.log 2>&1; echo "pytest exit $?" >> $O/a_pytest.log
tail -5 $O/a_pytest.log
for m in 1 16 64 128 256; do
  timeout 300 python tools/kbench.py gemm_int8 --m $m >> $O/a_kbench_streamk.log 2>&1
  timeout 300 python tools/kbench.py gemm_int8 --m $m --cluster >> $O/a_kbench_cluster.log 2>&1
done
cat $O/a_kbench_streamk.log | cut -c1-200; cat $O/a_kbench_cluster.log | cut -c1-200
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o $O/probe_tma_bw tools/probe_tma_bw.cu -lcuda > $O/a_probe_build.log 2>&1 && timeout 300 $O/probe_tma_bw > $O/a_probe.log 2>&1; tail -40 $O/a_probe.log
timeout 900 python bench.py --no-cpu-baseline > $O/a_bench.log 2> $O/a_bench.err; tail -n 1 $O/a_bench.log | cut -c1-1500; tail -3 $O/a_bench.err
