#!/bin/bash
# head gate at large batches: cfg3 (256 streams, 24 layers) and cfg5 (256 streams, 6 layers, generate_audio_batch)
mkdir -p gpurun_out
for v in 0 1; do
  PTTS_HEAD_GATE=$v PTTS_HEAD_GATE_MAX=4096 timeout 600 python bench.py --workload cfg3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('cfg3 gate $v:', d['value'], d['ms_per_step'], d['e2e']['value'])"
  PTTS_HEAD_GATE=$v PTTS_HEAD_GATE_MAX=4096 timeout 600 python bench.py --workload cfg5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('cfg5 gate $v:', d['value'], d['ms_per_step'], d['host_split'])"
done
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
