#!/bin/bash
# round-2 session ZF (run under gpurun --gpus N): AWQ kernel bench on cfg-4 shard shapes (N=1 part) and cfg 4 end to end
# at TP = N after the tiles x P grid rule
set -u
N=${1:-2}
O=gpurun_out; mkdir -p $O
if [ "$N" = "2" ]; then
  for tp in 1 2 4 8; do python tools/kbench.py awq --tp $tp --m 128 2>/dev/null; done > $O/zf_kbench_awq.jsonl
  grep "layer total" $O/zf_kbench_awq.jsonl
fi
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29513 bench.py --gpus $N --model qwen2.5-32b --quant awq --bs 128 --ctx 8192 --fixed-global-batch \
   --no-cpu-baseline --timeline-out $O/zf_timeline_cfg4_n$N.csv > $O/zf_bench_cfg4_n$N.log 2> $O/zf_bench_cfg4_n$N.err
echo "cfg4 exit $?"; grep '^{' $O/zf_bench_cfg4_n$N.log | tail -n1 | cut -c1-300; head -6 $O/zf_timeline_cfg4_n$N.csv | cut -c1-140
