#!/bin/bash
# round-2 GPU session K: elect.sync issue regions in every tcgen05 kernel + restructured two-tile extend kernel
set -u
O=gpurun_out; mkdir -p $O
for c in "300" "700,300 100,0"; do echo "== $c"; timeout 40 python tools/fa_probe.py $c 2>&1 | tail -3; done
timeout 900 python -m pytest tests -m gpu -q -x > $O/k_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/k_pytest_gpu.log; tail -6 $O/k_pytest_gpu.log
for poly in 0 2 4; do
  B200_FA_POLY=$poly timeout 60 python tools/kbench.py extend --ctx 8192 2>&1 | sed "s/^{/{\"poly\": $poly, /" | tee -a $O/k_kbench_extend_fa.jsonl
  B200_FA_POLY=$poly timeout 60 python tools/kbench.py extend --ctx 8192 --chunk 2048 2>&1 | sed "s/^{/{\"poly\": $poly, /" | tee -a $O/k_kbench_extend_fa.jsonl
done
B200_EXTEND_FA=0 timeout 60 python tools/kbench.py extend --ctx 8192 2>&1 | tee $O/k_kbench_extend_onetile.jsonl
timeout 120 python tools/kbench.py gemm_int8 --m 64 2>&1 | tee $O/k_kbench_int8.jsonl
timeout 120 python tools/kbench.py gemm_int8 --m 8192 2>&1 | tee -a $O/k_kbench_int8.jsonl
timeout 200 python tools/kbench.py awq --tp 1 2>&1 | cut -c1-200 | tee $O/k_kbench_awq.jsonl
timeout 200 python tools/kbench.py gemm_fp8 2>&1 | cut -c1-150 | tee $O/k_kbench_fp8_block.jsonl
timeout 300 python bench.py --no-cpu-baseline > $O/k_bench.log 2> $O/k_bench.err; tail -n1 $O/k_bench.log | cut -c1-330
B200_FA_POLY=0 timeout 300 ncu --set full --clock-control none --import-source on -k regex:extend_attn_fa_kernel -c 1 -o $O/k_extend_fa_poly0 -f \
   python tools/kbench.py extend --ctx 8192 > $O/k_ncu_poly0.log 2>&1; tail -1 $O/k_ncu_poly0.log
