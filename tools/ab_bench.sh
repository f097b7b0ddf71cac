#!/bin/bash
# A/B two builds of libnlbac_b200.so on the SAME box (clock / power-cap differences between boxes are +-4 %):
#   tools/ab_bench.sh ab_A.so ab_B.so   -> alternates A B A B, prints ms per step and the per-kernel split
P=$(ls -d neural*_b200)
for rep in 1 2; do
  for v in "$@"; do
    cp "$v" $P/libnlbac_b200.so
    python bench.py --steps 5 --warmup 3 --cpu-sample 64 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['kernel_ms_per_step']
print('$v', 'ms/step %.2f  fwd %.2f  passA %.2f  passB %.2f  sm_mhz %s' % (d['ms_per_step'], k['fwd_solve'], k['reverse_sweep'], k['dw_gemm'], d['clocks']['sm_mhz']))"
  done
done
