#!/bin/bash
# round-2 GPU session ZI: last regression of the round — GPU suite (-x, like the driver), smoke, headline bench line
set -u
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -x -q -m gpu > $O/zi_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/zi_pytest_gpu.log; tail -3 $O/zi_pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $O/zi_smoke.log 2>&1; echo "smoke exit $?" >> $O/zi_smoke.log; tail -2 $O/zi_smoke.log
timeout 600 python bench.py --no-cpu-baseline > $O/zi_bench.log 2> $O/zi_bench.err; echo "bench exit $?" >> $O/zi_bench.err; tail -n 1 $O/zi_bench.log | cut -c1-300
