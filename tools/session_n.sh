#!/bin/bash
# Multi-GPU bench exactly as the driver launches it: bash tools/session_n.sh N
set -u
N=$1; O=gpurun_out; mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 20 --warmup 3 > $O/n${N}_bench.log 2> $O/n${N}_bench.err; echo "exit $?" >> $O/n${N}_bench.err
grep '^{' $O/n${N}_bench.log | tail -1 | cut -c1-900; tail -3 $O/n${N}_bench.err
