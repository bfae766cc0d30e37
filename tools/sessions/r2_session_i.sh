#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 200 python -m pytest tests/test_attention_gpu.py tests/test_fullsize_gpu.py -m gpu -q -k "extend or chunked" > $O/i_pytest_extend.log 2>&1; echo "exit $?" >> $O/i_pytest_extend.log; tail -4 $O/i_pytest_extend.log
for poly in 0 2; do
B200_FA_POLY=$poly timeout 300 ncu --set full --clock-control none --import-source on -k regex:extend_attn_fa_kernel -c 1 -o $O/i_extend_fa_poly$poly -f \
   python tools/kbench.py extend --ctx 8192 > $O/i_ncu_poly$poly.log 2>&1; tail -2 $O/i_ncu_poly$poly.log
done
ls -la $O | grep ncu-rep
