#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 100 python tools/kbench.py decode --chunk 2048 2>&1 | tee $O/n_kbench_decode.jsonl
timeout 100 python tools/kbench.py decode --chunk 2048 --extra-splits 1 2>&1 | tee -a $O/n_kbench_decode.jsonl
timeout 100 python tools/kbench.py decode --chunk 2176 2>&1 | tee -a $O/n_kbench_decode.jsonl
timeout 100 python tools/kbench.py decode --ctx 4128 --chunk 2048 2>&1 | tee -a $O/n_kbench_decode.jsonl
timeout 100 python tools/kbench.py decode --ctx 4128 --chunk 2064 2>&1 | tee -a $O/n_kbench_decode.jsonl
timeout 600 python bench.py --workload prefill --quant fp8 --no-cpu-baseline > $O/n_bench_prefill_fp8.log 2> $O/n_bench_prefill_fp8.err; tail -n 1 $O/n_bench_prefill_fp8.log | cut -c1-300
timeout 600 python bench.py --workload prefill --quant int8 --no-cpu-baseline > $O/n_bench_prefill_int8.log 2> $O/n_bench_prefill_int8.err; tail -n 1 $O/n_bench_prefill_int8.log | cut -c1-300
