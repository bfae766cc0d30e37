#!/bin/bash
# round-2 multi-GPU session after the start barrier went back to relaxed flags and b200_tp_all_reduce was added:
#   bash tools/sessions/r2_session_mgpu2.sh N [cfg4] [cfg5]      (run under gpurun --gpus N)
# fused-exchange / all-reduce check, headline weak-scaling line, optionally config 4 (AWQ TP=N) and config 5 (70B TP=8).
set -u
N=${1:-2}
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 tools/tp_exchange_check.py > $O/n${N}_tp_exchange_check.jsonl 2> $O/n${N}_tp_exchange_check.err
echo "exchange check exit $?"; grep -c '"ok": true' $O/n${N}_tp_exchange_check.jsonl; grep '"ok": false' $O/n${N}_tp_exchange_check.jsonl | cut -c1-300
grep 'trace\|timing' $O/n${N}_tp_exchange_check.jsonl | cut -c1-400
timeout 600 $TR --master-port 29512 bench.py --gpus $N --no-cpu-baseline > $O/n${N}_bench.log 2> $O/n${N}_bench.err
echo "bench exit $?"; grep '^{' $O/n${N}_bench.log | tail -n1 | cut -c1-300
for what in "${@:2}"; do
  if [ "$what" = "cfg4" ]; then
    timeout 900 $TR --master-port 29513 bench.py --gpus $N --model qwen2.5-32b --quant awq --bs 128 --ctx 8192 --fixed-global-batch \
       --no-cpu-baseline > $O/n${N}_bench_cfg4.log 2> $O/n${N}_bench_cfg4.err
    echo "cfg4 exit $?"; grep '^{' $O/n${N}_bench_cfg4.log | tail -n1 | cut -c1-300
  fi
  if [ "$what" = "cfg5" ]; then
    timeout 1200 $TR --master-port 29514 bench.py --gpus $N --model llama-3.1-70b --bs 256 --ctx 32768 --shared-prefix 30720 --fixed-global-batch \
       --no-cpu-baseline > $O/n${N}_bench_cfg5.log 2> $O/n${N}_bench_cfg5.err
    echo "cfg5 exit $?"; grep '^{' $O/n${N}_bench_cfg5.log | tail -n1 | cut -c1-300
  fi
done
