#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "mimi or bench_shape or head_gated or long_context" -s 2>&1 | grep -v "^$" | tail -12
for v in 1 0; do
  PTTS_WIN_STREAM=$v timeout 300 python profiles/mimi_sweep.py 64 256 1024 2>&1 | grep -v WARN | sed "s/^/stream=$v /"
done
for v in 1 0 1 0; do
  PTTS_WIN_STREAM=$v timeout 300 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('stream=$v', d['value'], d['ms_per_step'], d['e2e']['value'])"
done
