#!/bin/bash
# Session K: whole GPU suite, smoke, bench (+reference arm), kernel benches, launch list, one ncu --set full capture of
# the fp8-KV decode kernel and of the shared-prefix pass.
set -u
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -x -q > $O/k_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/k_pytest_gpu.log; tail -4 $O/k_pytest_gpu.log
timeout 600 python __graft_entry__.py --smoke > $O/k_smoke.log 2>&1; echo "smoke exit $?" >> $O/k_smoke.log; tail -n 2 $O/k_smoke.log
timeout 900 python bench.py > $O/k_bench.log 2> $O/k_bench.err; echo "bench exit $?" >> $O/k_bench.err; tail -n 1 $O/k_bench.log | cut -c1-1200; tail -n 2 $O/k_bench.err
timeout 600 python bench.py --impl reference --steps 4 > $O/k_bench_ref.log 2> $O/k_bench_ref.err; tail -n 1 $O/k_bench_ref.log | cut -c1-600
(timeout 300 python tools/kbench.py decode --chunk 1024; timeout 300 python tools/kbench.py decode --fp8kv --chunk 1024; timeout 300 python tools/kbench.py decode --fp8kv --chunk 2048 --bs 128 --ctx 8192 --hq 20 --hkv 4; timeout 300 python tools/kbench.py decode --bs 256 --hq 8 --hkv 1 --ctx 32768 --shared 30720 --chunk 2048) > $O/k_kbench_attn.jsonl 2>&1; cat $O/k_kbench_attn.jsonl
timeout 600 python tools/profile_step.py --layers 2 --steps 3 > $O/k_profile_step.log 2>&1 && \
timeout 900 ncu --nvtx --nvtx-include "steps/" --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
   --log-file $O/k_launches.csv python tools/profile_step.py --layers 2 --steps 3 > $O/k_ncu_launch.log 2>&1
tail -n 2 $O/k_ncu_launch.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:decode_attn_fp8kv_kernel -c 1 -o $O/k_fp8kv_full \
   python tools/kbench.py decode --fp8kv --chunk 1024 > $O/k_ncu_fp8kv.log 2>&1; tail -n 2 $O/k_ncu_fp8kv.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:extend_attn_umma_kernel -c 1 -o $O/k_prefix_full \
   python tools/kbench.py decode --bs 256 --hq 8 --hkv 1 --ctx 32768 --shared 30720 --chunk 2048 > $O/k_ncu_prefix.log 2>&1; tail -n 2 $O/k_ncu_prefix.log
