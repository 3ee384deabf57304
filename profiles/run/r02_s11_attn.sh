#!/bin/bash
# window attention with batched projection-row loads + LDS/STS row epilogue: Mimi parity tests, Mimi sweep, bench
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "mimi or bench_shape or large_batch" 2>&1 | tail -4
timeout 300 python profiles/mimi_sweep.py 16 64 256 1024 2>&1 | grep -v WARNING | tail -4
timeout 300 python bench.py > gpurun_out/r02_s11_bench_attn.json 2> gpurun_out/r02_s11_bench_attn.err; echo "bench exit $?"
python - <<'P'
import json
d = json.loads(open("gpurun_out/r02_s11_bench_attn.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["clocks"], d["kernel_ms_share"])
P
