#!/bin/bash
# row-major persistent GEMM: pipeline depth vs shared-memory budget on the Mimi pass shapes
for kb in 112 144 176 208; do
  echo "== PTTS_TC_ROWS_SMEM_KB=$kb"
  PTTS_TC_ROWS_SMEM_KB=$kb timeout 120 ./profiles/probes/gemm_rows_bench.bin 2>&1 | grep -v "no act\|N1536\|N1024\|N4096"
done
