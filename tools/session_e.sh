#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_attention_gpu.py -x -q -k "int8_out" > $O/e_pytest.log 2>&1; tail -3 $O/e_pytest.log
timeout 900 python bench.py --no-cpu-baseline > $O/e_bench.log 2> $O/e_bench.err; tail -n 1 $O/e_bench.log | cut -c1-120; tail -n 1 $O/e_bench.log | grep -o '"roofline.*' | cut -c1-600; tail -3 $O/e_bench.err
B200_FUSE_ATTN_QUANT=0 timeout 900 python bench.py --no-cpu-baseline > $O/e_bench_nofuse.log 2> $O/e_bench_nofuse.err; tail -n 1 $O/e_bench_nofuse.log | cut -c1-120;  tail -n 1 $O/e_bench_nofuse.log | grep -o '"roofline.*' | cut -c1-600
