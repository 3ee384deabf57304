#!/bin/bash
# final code (uniform group schedule): suite, smoke, driver-shape bench, default bench, cfg3, cfg5
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_final.log 2>&1; echo "suite exit $?" >> gpurun_out/r02_pytest_gpu_final.log; tail -3 gpurun_out/r02_pytest_gpu_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_final.log 2>&1; tail -2 gpurun_out/r02_smoke_final.log
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_driver_shape.json 2>/dev/null
python bench.py > gpurun_out/r02_bench_final.json 2> gpurun_out/r02_bench_final.err
python bench.py --workload cfg3 --steps 48 > gpurun_out/r02_bench_cfg3_final.json 2>/dev/null
python bench.py --workload cfg5 > gpurun_out/r02_bench_cfg5_final.json 2>/dev/null
python - <<'P'
import json
for f in ("r02_bench_driver_shape","r02_bench_final","r02_bench_cfg3_final","r02_bench_cfg5_final"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["value"], d.get("ms_per_step"), d.get("e2e", {}).get("value"), d.get("first_chunk"), (d.get("roofline") or {}).get("frac"), d.get("step_roofline"), (d.get("roofline_in_graph") or {}).get("frac"))
    except Exception as e:
        print(f, "ERR", e)
P
