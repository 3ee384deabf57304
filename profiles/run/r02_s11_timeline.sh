#!/bin/bash
mkdir -p gpurun_out
for mode in "PTTS_X=gated" "PTTS_HEAD_GATE=0" "PTTS_PROFILE_PART=3"; do
  echo "== $mode"; env $mode timeout 300 python profiles/chain_timeline.py 2>/dev/null | grep -v WARN
done
