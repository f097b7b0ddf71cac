#!/bin/bash
# round-2 GPU call 1: torchdiffeq probe, tensor-pipe peaks, A/B of the TC kernel variants (ab/*.so)
P=$(ls -d neural*_b200)
mkdir -p gpurun_out
{ python -c "import torchdiffeq; print('torchdiffeq', torchdiffeq.__version__)"; echo "exit $?";
  timeout 30 python -m pip download --no-deps -d /tmp/td torchdiffeq==0.2.3; echo "pip exit $?"; } > gpurun_out/probe_torchdiffeq.txt 2>&1
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.txt; nproc >> gpurun_out/smi.txt
cp ab/v3.so $P/libnlbac_b200.so
timeout 120 python tools/umma_peak.py 3 > gpurun_out/umma_peak.json 2> gpurun_out/umma_peak.err
for v in base v1 v2 v3 v4 v5; do
  cp ab/$v.so $P/libnlbac_b200.so
  echo "== $v" >> gpurun_out/ab_check.txt
  timeout 180 python tools/ab_check.py >> gpurun_out/ab_check.txt 2>&1 || echo "FAILED $v $?" >> gpurun_out/ab_check.txt
done
for rep in 1 2; do
  for v in base v1 v2 v3 v4 v5; do
    cp ab/$v.so $P/libnlbac_b200.so
    timeout 180 python bench.py --steps 5 --warmup 3 --cpu-sample 64 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['kernel_ms_per_step']
print('$v', 'ms/step %.2f  fwd %.2f  sweep %.2f  dw %.2f  sm_mhz %s' % (d['ms_per_step'], k['fwd_solve'], k['reverse_sweep'], k['dw_gemm'], d['clocks']['sm_mhz']))" >> gpurun_out/ab_bench.txt 2>&1
  done
done
cat gpurun_out/probe_torchdiffeq.txt gpurun_out/umma_peak.json gpurun_out/ab_check.txt gpurun_out/ab_bench.txt
