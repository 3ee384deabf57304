#!/bin/bash
# Evidence run of the final code (after the streaming window attention): suite, smoke, benches, sweep, launch list of a 32-step bench
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_final.log 2>&1; echo "suite exit $?" >> gpurun_out/r02_pytest_gpu_final.log; tail -3 gpurun_out/r02_pytest_gpu_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_final.log 2>&1; tail -2 gpurun_out/r02_smoke_final.log
python bench.py > gpurun_out/r02_bench_final.json 2> gpurun_out/r02_bench_final.err
python bench.py --impl reference --steps 24 --warmup 4 > gpurun_out/r02_bench_reference_final.json 2>/dev/null
python bench.py --workload cfg3 --steps 48 > gpurun_out/r02_bench_cfg3_final.json 2>/dev/null
python bench.py --workload cfg5 > gpurun_out/r02_bench_cfg5_final.json 2>/dev/null
python bench.py --workload cfg5 > gpurun_out/r02_bench_cfg5_final2.json 2>/dev/null
python profiles/mimi_sweep.py 2>&1 | grep -v WARN > gpurun_out/r02_mimi_sweep_cfg4_final.txt
python - <<'P'
import json
for f in ("r02_bench_final","r02_bench_reference_final","r02_bench_cfg3_final","r02_bench_cfg5_final","r02_bench_cfg5_final2"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["value"], d.get("ms_per_step"), d.get("e2e", {}).get("value"), d.get("first_chunk"), (d.get("roofline") or {}).get("frac"), d.get("step_roofline"))
    except Exception as e:
        print(f, "ERR", e)
P
tail -6 gpurun_out/r02_mimi_sweep_cfg4_final.txt
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file gpurun_out/r02_launches_bench.csv python bench.py --steps 32 --warmup 4 > gpurun_out/r02_ncu_launches.log 2>&1
echo "ncu launch list exit $?"; wc -l gpurun_out/r02_launches_bench.csv
