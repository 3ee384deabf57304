#!/bin/bash
# same-box A/B: skinny GEMM at 112 registers (default build) vs 168 (build/ab), alternating, default bench
mkdir -p gpurun_out
for i in 1 2 3; do
  for v in 112 168; do
    if [ $v = 168 ]; then export PTTS_NATIVE_LIB=$PWD/pocket_tts_b200/build/ab/libptts_regs168.so; else unset PTTS_NATIVE_LIB; fi
    timeout 300 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('regs $v run $i:', d['value'], d['ms_per_step'], d['e2e']['value'])"
  done
done
unset PTTS_NATIVE_LIB
for v in 112 168; do
  if [ $v = 168 ]; then export PTTS_NATIVE_LIB=$PWD/pocket_tts_b200/build/ab/libptts_regs168.so; else unset PTTS_NATIVE_LIB; fi
  PTTS_PROFILE_PART=3 timeout 300 python bench.py 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('regs $v chain only:', d['ms_per_step'])"
done
