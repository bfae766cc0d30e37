#!/bin/bash
# round-2 multi-GPU session: bash tools/sessions/r2_session_mgpu.sh N   (N = 2, 4 or 8; run under gpurun --gpus N)
# headline weak-scaling line + kernel timeline, the fused-exchange check, config 4 (AWQ TP=N) and, at N=8, config 5.
set -u
N=${1:-2}
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi --query-gpu=index,name,clocks.sm --format=csv > $O/m${N}_smi.txt 2>&1
timeout 600 $TR --master-port 29511 tools/tp_exchange_check.py > $O/m${N}_tp_exchange_check.jsonl 2> $O/m${N}_tp_exchange_check.err
echo "exchange check exit $?"; grep -c '"ok": true' $O/m${N}_tp_exchange_check.jsonl; grep '"ok": false' $O/m${N}_tp_exchange_check.jsonl | cut -c1-300
timeout 600 $TR --master-port 29512 bench.py --gpus $N --timeline-out $O/m${N}_timeline_headline.csv > $O/m${N}_bench.log 2> $O/m${N}_bench.err
echo "bench exit $?"; grep '^{' $O/m${N}_bench.log | tail -n1 | cut -c1-400
head -14 $O/m${N}_timeline_headline.csv | cut -c1-150
# same step with programmatic dependent launch off: kernels no longer start under their predecessor's tail, so the
# CUPTI durations are each kernel's own time (the step itself is slower; attribution only, not a bench value)
B200_PDL=0 timeout 600 $TR --master-port 29515 bench.py --gpus $N --steps 5 --no-cpu-baseline --timeline-out $O/m${N}_timeline_headline_nopdl.csv > $O/m${N}_bench_nopdl.log 2>&1
grep '^{' $O/m${N}_bench_nopdl.log | tail -n1 | cut -c1-260
head -14 $O/m${N}_timeline_headline_nopdl.csv | cut -c1-150
timeout 900 $TR --master-port 29513 bench.py --gpus $N --model qwen2.5-32b --quant awq --bs 128 --ctx 8192 --fixed-global-batch \
   --timeline-out $O/m${N}_timeline_cfg4.csv > $O/m${N}_bench_cfg4.log 2> $O/m${N}_bench_cfg4.err
echo "cfg4 exit $?"; grep '^{' $O/m${N}_bench_cfg4.log | tail -n1 | cut -c1-400; tail -n 2 $O/m${N}_bench_cfg4.err | cut -c1-300
head -12 $O/m${N}_timeline_cfg4.csv | cut -c1-150
if [ "$N" = "8" ]; then
  timeout 1200 $TR --master-port 29514 bench.py --gpus 8 --model llama-3.1-70b --bs 256 --ctx 32768 --shared-prefix 30720 --fixed-global-batch \
     --timeline-out $O/m8_timeline_cfg5.csv > $O/m8_bench_cfg5.log 2> $O/m8_bench_cfg5.err
  echo "cfg5 exit $?"; grep '^{' $O/m8_bench_cfg5.log | tail -n1 | cut -c1-600; tail -n 2 $O/m8_bench_cfg5.err | cut -c1-300
  head -12 $O/m8_timeline_cfg5.csv | cut -c1-150
fi
