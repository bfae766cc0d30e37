#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
python tools/ncu_gemm.py 64 > $O/b_plain.log 2>&1 || { tail -5 $O/b_plain.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_streamk --launch-skip 4 -c 4 -o $O/streamk_m64 -f python tools/ncu_gemm.py 64 > $O/b_ncu1.log 2>&1; tail -2 $O/b_ncu1.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_streamk --launch-skip 4 -c 4 -o $O/streamk_m16 -f python tools/ncu_gemm.py 16 > $O/b_ncu2.log 2>&1; tail -2 $O/b_ncu2.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_umma --launch-skip 4 -c 4 -o $O/cluster_m64 -f python tools/ncu_gemm.py 64 cluster > $O/b_ncu3.log 2>&1; tail -2 $O/b_ncu3.log
ls -la $O/*.ncu-rep
