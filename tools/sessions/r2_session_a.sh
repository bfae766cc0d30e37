#!/bin/bash
# round-2 GPU session A: parity (new per-row asserts, reference-function goldens, full-size cfg-3 cases) + fp8 GEMM baselines
set -u
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/a_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_fullsize_gpu.py > $O/a_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/a_pytest_gpu.log
tail -25 $O/a_pytest_gpu.log
timeout 900 python -m pytest tests/test_fullsize_gpu.py -m gpu -q > $O/a_pytest_fullsize.log 2>&1; echo "pytest exit $?" >> $O/a_pytest_fullsize.log
tail -40 $O/a_pytest_fullsize.log
timeout 300 python __graft_entry__.py --smoke > $O/a_smoke.log 2>&1; echo "smoke exit $?" >> $O/a_smoke.log; tail -3 $O/a_smoke.log
timeout 600 python tools/kbench.py gemm_fp8 > $O/a_kbench_fp8_block.jsonl 2> $O/a_kbench_fp8_block.err; tail -20 $O/a_kbench_fp8_block.jsonl; tail -3 $O/a_kbench_fp8_block.err
timeout 600 python tools/kbench.py gemm_fp8 --chan > $O/a_kbench_fp8_chan.jsonl 2> $O/a_kbench_fp8_chan.err; tail -20 $O/a_kbench_fp8_chan.jsonl
