#!/bin/bash
# skinny GEMM capped at 112 registers (co-residency with the Mimi row-major kernel): parity + bench, twice
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -4
for i in 1 2; do
timeout 300 python bench.py > gpurun_out/r02_s11_bench_regs$i.json 2> gpurun_out/r02_s11_bench_regs$i.err; echo "bench exit $?"
done
python - <<'P'
import json
for i in (1, 2):
    d = json.loads(open(f"gpurun_out/r02_s11_bench_regs{i}.json").read().strip().splitlines()[-1])
    print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["first_chunk"], d["kernel_ms_share"])
P
PTTS_PROFILE_PART=3 timeout 300 python bench.py --steps 96 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('chain only', d['ms_per_step'])"
