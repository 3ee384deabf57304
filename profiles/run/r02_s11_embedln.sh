#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "layernorm_folded or bench_shape or head_gated or pipelined" 2>&1 | tail -3
for v in 1 0 1 0; do
  PTTS_EMBED_LN=$v timeout 300 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('embed_ln=$v', d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'])"
done
