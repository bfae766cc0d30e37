#!/bin/bash
# Session M (round-end evidence): whole GPU suite, smoke, bench (+ reference arm), kernel benches, launch list.
# Every command runs under its own timeout.
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/m_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/m_pytest_gpu.log; tail -3 $O/m_pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $O/m_smoke.log 2>&1; echo "smoke exit $?" >> $O/m_smoke.log; tail -n 2 $O/m_smoke.log
timeout 600 python bench.py > $O/m_bench.log 2> $O/m_bench.err; echo "bench exit $?" >> $O/m_bench.err; tail -n 1 $O/m_bench.log | cut -c1-600; tail -n 1 $O/m_bench.err
timeout 400 python bench.py --impl reference --steps 4 > $O/m_bench_ref.log 2> $O/m_bench_ref.err; tail -n 1 $O/m_bench_ref.log | cut -c1-300
(timeout 120 python tools/kbench.py gemm_int8; timeout 120 python tools/kbench.py awq) > $O/m_kbench_gemm.jsonl 2>&1; tail -n 30 $O/m_kbench_gemm.jsonl | cut -c1-220
timeout 300 python tools/profile_step.py --layers 2 --steps 3 > $O/m_profile_step.log 2>&1 && \
timeout 600 ncu --nvtx --nvtx-include "steps/" --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
   --log-file $O/m_launches.csv python tools/profile_step.py --layers 2 --steps 3 > $O/m_ncu_launch.log 2>&1
tail -n 2 $O/m_ncu_launch.log
