#!/bin/bash
# 8-GPU runs of config 5: weak (2^20 per GPU), strong 2^23 and 2^26 total (Unicycle), strong 2^23 (PVTOL)
N=${1:-8}; out=gpurun_out/r2_scale_n$N.jsonl; : > $out; port=29600
run() { port=$((port+1)); python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --steps 3 --warmup 2 --sweep-only "$@" 2>/dev/null >> $out || echo "{\"failed\": \"$*\"}" >> $out; }
run
run --scaling strong --log2m-total 23
run --scaling strong --log2m-total 26
run --scaling strong --log2m-total 23 --env P
python - <<PY
import json
for l in open('$out'):
    d=json.loads(l)
    if 'value' in d: print(d['n_gpus'], d['scaling'], d['config']['workload'][:70], 'value %.3e' % d['value'], 'ms %.1f' % d['ms_per_step'])
    else: print(d)
PY
