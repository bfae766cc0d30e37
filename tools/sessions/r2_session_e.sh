#!/bin/bash
# round-2 GPU session E: full state check after the container restore — parity suite, smoke, headline + prefill bench
# lines, kernel benches of every GEMM family, launch list of the decode step.
set -u
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/e_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q > $O/e_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/e_pytest_gpu.log
tail -12 $O/e_pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $O/e_smoke.log 2>&1; echo "smoke exit $?" >> $O/e_smoke.log; tail -2 $O/e_smoke.log
timeout 600 python bench.py > $O/e_bench.log 2> $O/e_bench.err; echo "bench exit $?" >> $O/e_bench.err; tail -n 1 $O/e_bench.log | cut -c1-1500; tail -n 2 $O/e_bench.err
timeout 600 python bench.py --workload prefill --quant fp8 > $O/e_bench_prefill_fp8.log 2> $O/e_bench_prefill_fp8.err; echo "exit $?" >> $O/e_bench_prefill_fp8.err; tail -n 1 $O/e_bench_prefill_fp8.log | cut -c1-1500; tail -n 2 $O/e_bench_prefill_fp8.err
timeout 600 python bench.py --workload prefill --quant int8 --no-cpu-baseline > $O/e_bench_prefill_int8.log 2> $O/e_bench_prefill_int8.err; echo "exit $?" >> $O/e_bench_prefill_int8.err; tail -n 1 $O/e_bench_prefill_int8.log | cut -c1-1500
timeout 300 python tools/kbench.py gemm_fp8 > $O/e_kbench_fp8_block.jsonl 2> $O/e_kbench_fp8_block.err; cut -c1-230 $O/e_kbench_fp8_block.jsonl
timeout 300 python tools/kbench.py gemm_int8 --m 64 > $O/e_kbench_int8.jsonl 2>&1
timeout 300 python tools/kbench.py gemm_int8 --m 8192 >> $O/e_kbench_int8.jsonl 2>&1; cat $O/e_kbench_int8.jsonl
timeout 300 python tools/kbench.py awq --tp 1 > $O/e_kbench_awq.jsonl 2>&1; cut -c1-200 $O/e_kbench_awq.jsonl
timeout 300 python tools/kbench.py extend --ctx 8192 > $O/e_kbench_extend.jsonl 2>&1
timeout 300 python tools/kbench.py extend --ctx 8192 --chunk 2048 >> $O/e_kbench_extend.jsonl 2>&1; cat $O/e_kbench_extend.jsonl
timeout 600 python tools/profile_step.py --layers 2 --steps 3 > $O/e_profile_step.log 2>&1 && \
timeout 900 ncu --nvtx --nvtx-include "steps/" --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
   --log-file $O/e_launches.csv python tools/profile_step.py --layers 2 --steps 3 > $O/e_ncu_launch.log 2>&1
tail -n 2 $O/e_ncu_launch.log
