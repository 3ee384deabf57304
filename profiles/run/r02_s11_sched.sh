#!/bin/bash
# group schedule of a blocking call (PTTS_SCHED): the driver's bench shape (20 steps after 5 warm-up) and the default (96 / 4)
for p in 0 1 2 3; do
  PTTS_SCHED=$p timeout 300 python bench.py --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('sched $p  steps 20:', d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'])"
done
for p in 0 2 3; do
  PTTS_SCHED=$p timeout 300 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('sched $p  steps 96:', d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['first_chunk'])"
done
for p in 0 2; do
  PTTS_SCHED=$p timeout 300 python bench.py --workload cfg5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('sched $p  cfg5:', d['value'], d['ms_per_step'], d['host_split'])"
done
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "pipelined or head_gated or bench_shape or long_context" 2>&1 | tail -2
