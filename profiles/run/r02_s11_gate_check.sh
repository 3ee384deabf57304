#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "head_gated or pipelined or bench_shape" 2>&1 | tail -4
for w in cfg2 cfg5 cfg3; do
  timeout 600 python bench.py --workload $w 2>/dev/null > gpurun_out/r02_s11_gate_$w.json
  python -c "
import json
d=json.loads(open('gpurun_out/r02_s11_gate_$w.json').read().strip().splitlines()[-1]); print('$w:', d['value'], d['ms_per_step'], d['e2e']['value'], d.get('first_chunk'))"
done
