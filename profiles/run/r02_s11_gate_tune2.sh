#!/bin/bash
mkdir -p gpurun_out
run() { echo -n "$1: "; env $1 PTTS_HEAD_GATE=1 timeout 300 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'])"; }
run "PTTS_X=base"
run "PTTS_GATE_EARLY=2"
run "PTTS_X=base"
run "PTTS_GATE_EARLY=2"
run "PTTS_GATE_EARLY=2 PTTS_GATE_RESERVE=16"
run "PTTS_GATE_EARLY=2 PTTS_FUSED_SMS=-1"
run "PTTS_FUSED_SMS=-1"
