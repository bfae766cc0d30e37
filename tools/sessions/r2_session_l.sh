#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
for c in "300" "700,300 100,0" "333,200 16,137 1"; do echo "== $c"; timeout 40 python tools/fa_probe.py $c 2>&1 | tail -3 | cut -c1-120; done
timeout 200 python -m pytest tests/test_attention_gpu.py tests/test_fullsize_gpu.py -m gpu -q -k "extend or chunked" > $O/l_pytest_extend.log 2>&1; echo "exit $?" >> $O/l_pytest_extend.log; tail -3 $O/l_pytest_extend.log
for poly in 0 2 4; do
  B200_FA_POLY=$poly timeout 60 python tools/kbench.py extend --ctx 8192 2>&1 | sed "s/^{/{\"poly\": $poly, /" | tee -a $O/l_kbench_extend_fa.jsonl
  B200_FA_POLY=$poly timeout 60 python tools/kbench.py extend --ctx 8192 --chunk 2048 2>&1 | sed "s/^{/{\"poly\": $poly, /" | tail -1 | tee -a $O/l_kbench_extend_fa.jsonl
done
B200_FA_POLY=${1:-0} timeout 300 ncu --set full --clock-control none --import-source on -k regex:extend_attn_fa_kernel -c 1 -o $O/l_extend_fa -f \
   python tools/kbench.py extend --ctx 8192 > $O/l_ncu.log 2>&1; tail -1 $O/l_ncu.log
