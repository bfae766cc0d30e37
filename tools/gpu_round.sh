#!/bin/bash
# One GPU-box session.  Usage: bash tools/gpu_round.sh [steps...]   steps: test smoke bench ab ref kbench launches full
# Everything lands in gpurun_out/ (scratch); tools/summarize_profiles.py turns it into profiles/.
set -u
O=gpurun_out
mkdir -p $O
STEPS="${*:-test smoke bench ref kbench launches full}"
has() { [[ " $STEPS " == *" $1 "* ]]; }
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
if has test; then
  timeout 2400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/pytest_gpu.log
  tail -15 $O/pytest_gpu.log
fi
if has smoke; then
  timeout 600 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke exit $?" >> $O/smoke.log
  tail -n 3 $O/smoke.log
fi
if has bench; then
  timeout 900 python bench.py > $O/bench.log 2> $O/bench.err; echo "bench exit $?" >> $O/bench.err
  tail -n 2 $O/bench.log; tail -n 3 $O/bench.err
fi
if has ab; then
  B200_PDL=0 timeout 600 python bench.py --no-cpu-baseline > $O/bench_nopdl.log 2>&1; tail -n 1 $O/bench_nopdl.log | cut -c1-400
  B200_FUSE_STEPS=0 timeout 600 python bench.py --no-cpu-baseline > $O/bench_nofuse.log 2>&1; tail -n 1 $O/bench_nofuse.log | cut -c1-400
fi
if has ref; then
  timeout 600 python bench.py --impl reference --steps 4 > $O/bench_ref.log 2> $O/bench_ref.err
  tail -n 1 $O/bench_ref.log
fi
if has kbench; then
  timeout 600 python tools/kbench.py decode > $O/kbench_decode.log 2>&1
  timeout 600 python tools/kbench.py gemm_int8 > $O/kbench_int8.log 2>&1
  tail -n 40 $O/kbench_decode.log $O/kbench_int8.log
fi
if has launches; then
  timeout 600 python tools/profile_step.py --layers 2 --steps 3 > $O/profile_step.log 2>&1 && \
  timeout 900 ncu --nvtx --nvtx-include "steps/" --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
     --log-file $O/launches.csv python tools/profile_step.py --layers 2 --steps 3 > $O/ncu_launch.log 2>&1
  tail -n 3 $O/ncu_launch.log
fi
if has full; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:decode_attn_kernel -c 2 -o $O/decode_attn_full -f \
     python tools/profile_step.py --layers 2 --steps 2 > $O/ncu_full.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_umma_kernel -c 4 -o $O/gemm_decode_full -f \
     python tools/profile_step.py --layers 1 --steps 1 > $O/ncu_full_gemm.log 2>&1
fi
ls -la $O | head -40
