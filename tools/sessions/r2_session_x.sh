#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
show() { python - "$1" <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1]); r=d['roofline']
print(sys.argv[1], round(d['value']), round(d['ms_per_step'],2), [(g['k'],g['n'],round(g['ms'],3),round(g['tflops'])) for g in r['all_gemms']], 'attn', round(r['extend_attention']['ms_per_launch'],3), d['clocks']['sm_mhz'])
PY
}
B200_GEMM_PAIR=1 timeout 300 python bench.py --workload prefill --quant int8 --no-cpu-baseline --steps 10 > $O/x_pair1.log 2>&1; show $O/x_pair1.log
B200_GEMM_PAIR=0 timeout 300 python bench.py --workload prefill --quant int8 --no-cpu-baseline --steps 10 > $O/x_pair0.log 2>&1; show $O/x_pair0.log
B200_GEMM_PAIR=1 B200_PDL=0 timeout 300 python bench.py --workload prefill --quant int8 --no-cpu-baseline --steps 10 > $O/x_pair1_nopdl.log 2>&1; show $O/x_pair1_nopdl.log
