#!/bin/bash
# head-gate tuning on one box: segment -> frame-slot maps, SMs reserved for the head, early gate
mkdir -p gpurun_out
run() { echo -n "$1: "; env $1 PTTS_HEAD_GATE=1 timeout 300 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'])"; }
run "PTTS_X=base"
run "PTTS_GATE_MAP=-1,0,1,2"
run "PTTS_GATE_MAP=0,0,1,2"
run "PTTS_GATE_MAP=-1,-1,0,1"
run "PTTS_GATE_MAP=0,1,2,2"
run "PTTS_GATE_RESERVE=0"
run "PTTS_GATE_RESERVE=16"
run "PTTS_GATE_EARLY=1"
run "PTTS_X=base"
PTTS_HEAD_GATE=0 timeout 300 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ungated', d['value'], d['ms_per_step'], d['e2e']['value'])"
