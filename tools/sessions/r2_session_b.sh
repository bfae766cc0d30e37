#!/bin/bash
# round-2 GPU session B: new block-FP8 kernels (stream-K decode + persistent prefill): parity, then timing
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_fp8_gpu.py tests/test_quant_golden_gpu.py -m gpu -q -x > $O/b_pytest_fp8.log 2>&1; echo "pytest exit $?" >> $O/b_pytest_fp8.log
tail -25 $O/b_pytest_fp8.log
timeout 600 python -m pytest tests/test_fullsize_gpu.py -m gpu -q -k block > $O/b_pytest_fullsize.log 2>&1; echo "pytest exit $?" >> $O/b_pytest_fullsize.log
tail -25 $O/b_pytest_fullsize.log
timeout 600 python tools/kbench.py gemm_fp8 > $O/b_kbench_fp8_block.jsonl 2> $O/b_kbench_fp8_block.err; cut -c1-260 $O/b_kbench_fp8_block.jsonl; tail -3 $O/b_kbench_fp8_block.err
