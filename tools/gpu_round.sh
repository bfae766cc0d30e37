#!/bin/bash
# One GPU-box session: parity tests, smoke, bench, kernel micro-benchmarks, ncu launch list + full capture.
# Everything lands in gpurun_out/ (scratch); summaries are copied into profiles/ by hand afterwards.
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
lscpu | head -30 > $O/host_cpu.txt 2>&1
grep -o -w -E 'avx512f|avx512_bf16|amx_bf16|amx_int8|avx512_vnni' /proc/cpuinfo | sort | uniq -c >> $O/host_cpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
timeout 600 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke exit $?" >> $O/smoke.log
tail -3 $O/smoke.log
timeout 900 python bench.py > $O/bench.log 2> $O/bench.err; echo "bench exit $?" >> $O/bench.err
tail -2 $O/bench.log; tail -3 $O/bench.err
timeout 600 python bench.py --impl reference --steps 4 > $O/bench_ref.log 2> $O/bench_ref.err
tail -1 $O/bench_ref.log
timeout 600 python tools/kbench.py decode > $O/kbench_decode.log 2>&1
timeout 600 python tools/kbench.py gemm_int8 > $O/kbench_int8.log 2>&1
tail -30 $O/kbench_decode.log $O/kbench_int8.log
timeout 600 python tools/profile_step.py --layers 2 --steps 3 > $O/profile_step.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
   python tools/profile_step.py --layers 2 --steps 3 > $O/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:decode_attn_kernel -c 2 -o $O/decode_attn_full -f \
   python tools/profile_step.py --layers 2 --steps 2 > $O/ncu_full.log 2>&1
ls -la $O
