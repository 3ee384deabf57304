#!/bin/bash
# head-gated Mimi passes (PTTS_HEAD_GATE=1): parity of the pipelined loop, then A/B of the default bench on one box
mkdir -p gpurun_out
PTTS_HEAD_GATE=1 timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "pipelined or bench_shape or long_context or large_batch" 2>&1 | tail -4
for i in 1 2; do
  for v in 0 1; do
    PTTS_HEAD_GATE=$v timeout 300 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('gate $v run $i:', d['value'], d['ms_per_step'], d['e2e']['value'], d['first_chunk']['p50_ms_batch_64'])"
  done
done
