#!/bin/bash
# round-2 GPU session W: full regression + bench lines with the pair GEMM / two-tile extend / FP8-KV shared prefix
set -u
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q > $O/w_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/w_pytest_gpu.log; tail -6 $O/w_pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $O/w_smoke.log 2>&1; echo "smoke exit $?" >> $O/w_smoke.log; tail -2 $O/w_smoke.log
timeout 600 python bench.py > $O/w_bench.log 2> $O/w_bench.err; echo "bench exit $?" >> $O/w_bench.err; tail -n 1 $O/w_bench.log | cut -c1-400
timeout 600 python bench.py --workload prefill --quant int8 --no-cpu-baseline > $O/w_bench_prefill_int8.log 2> $O/w_bench_prefill_int8.err; tail -n 1 $O/w_bench_prefill_int8.log | cut -c1-300
timeout 600 python bench.py --workload prefill --quant fp8 > $O/w_bench_prefill_fp8.log 2> $O/w_bench_prefill_fp8.err; tail -n 1 $O/w_bench_prefill_fp8.log | cut -c1-300
timeout 600 python bench.py --impl reference --steps 4 > $O/w_bench_ref.log 2> $O/w_bench_ref.err; tail -n 1 $O/w_bench_ref.log | cut -c1-300
timeout 600 python tools/profile_step.py --layers 2 --steps 3 > $O/w_profile_step.log 2>&1 && \
timeout 900 ncu --nvtx --nvtx-include "steps/" --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
   --log-file $O/w_launches.csv python tools/profile_step.py --layers 2 --steps 3 > $O/w_ncu_launch.log 2>&1
tail -n 1 $O/w_ncu_launch.log
