#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 --steps 20 --warmup 3; }
timeout 600 bash -c "$(declare -f run); run 29521" > $O/g_bench_n2_fused.log 2> $O/g_bench_n2_fused.err; grep '^{' $O/g_bench_n2_fused.log | tail -1 | cut -c1-1500; tail -2 $O/g_bench_n2_fused.err
B200_TP_FUSED_EXCHANGE=0 timeout 600 bash -c "$(declare -f run); run 29522" > $O/g_bench_n2_nccl.log 2> $O/g_bench_n2_nccl.err; grep '^{' $O/g_bench_n2_nccl.log | tail -1 | cut -c1-700
