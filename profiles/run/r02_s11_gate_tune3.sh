#!/bin/bash
run() { echo -n "$1: "; env $1 timeout 300 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'])"; }
run "PTTS_X=base"
run "PTTS_GATE_SEG0_SMS=74"
run "PTTS_GATE_SEG0_SMS=110"
run "PTTS_MIMI_GROUP=8"
run "PTTS_X=base"
