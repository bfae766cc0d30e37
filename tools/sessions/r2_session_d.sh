#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_fp8_gpu.py tests/test_quant_golden_gpu.py -m gpu -q -x > $O/d_pytest_fp8.log 2>&1; echo "pytest exit $?" >> $O/d_pytest_fp8.log
tail -5 $O/d_pytest_fp8.log
timeout 600 python -m pytest tests/test_fullsize_gpu.py -m gpu -q -k block > $O/d_pytest_fullsize.log 2>&1; tail -3 $O/d_pytest_fullsize.log
timeout 300 python tools/kbench.py gemm_fp8 2>&1 | cut -c1-200 | tee $O/d_kbench.log
timeout 300 python tools/gemm_trace.py 64 block 2>&1 | grep '"block"' | cut -c1-600
