#!/bin/bash
# config-5 sweep on ONE GPU: M = 2^20 .. 2^26 Unicycle, 2^20 / 2^24 PVTOL (TF32 tier headline + FP32 / 3xTF32 tiers at 2^20 only)
out=gpurun_out/r2_sweep.jsonl; : > $out
for lm in 20 22 24 26; do
  st=3; [ $lm -ge 24 ] && st=2
  python bench.py --steps $st --warmup 1 --log2m $lm --cpu-sample 0 --sweep-only 2>/dev/null >> $out || echo "{\"failed\": $lm}" >> $out
done
for lm in 20 24; do
  python bench.py --steps 2 --warmup 1 --log2m $lm --env P --cpu-sample 0 --sweep-only 2>/dev/null >> $out || echo "{\"failed_P\": $lm}" >> $out
done
python - <<'PY'
import json
for l in open('gpurun_out/r2_sweep.jsonl'):
    d=json.loads(l)
    if 'value' in d: print(d['config']['workload'][:60], 'value %.3e' % d['value'], 'ms %.1f' % d['ms_per_step'], d['kernel_ms_per_step'])
    else: print(d)
PY
