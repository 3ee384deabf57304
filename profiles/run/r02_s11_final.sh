#!/bin/bash
# Evidence run of the final code: whole GPU suite, smoke, default bench (twice), reference arm, cfg3, cfg5, cfg4 sweep, Mimi
# kernel times, ncu launch list of the default bench command, ncu --set full of the top kernels.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu_final.log 2>&1; echo "suite exit $?" >> gpurun_out/r02_pytest_gpu_final.log; tail -3 gpurun_out/r02_pytest_gpu_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke_final.log 2>&1; tail -2 gpurun_out/r02_smoke_final.log
python bench.py > gpurun_out/r02_bench_final.json 2> gpurun_out/r02_bench_final.err
python bench.py > gpurun_out/r02_bench_final2.json 2>/dev/null
python bench.py --impl reference --steps 24 --warmup 4 > gpurun_out/r02_bench_reference_final.json 2>/dev/null
python bench.py --workload cfg3 --steps 48 > gpurun_out/r02_bench_cfg3_final.json 2>/dev/null
python bench.py --workload cfg5 > gpurun_out/r02_bench_cfg5_final.json 2>/dev/null
python profiles/mimi_sweep.py 2>&1 | grep -v WARN > gpurun_out/r02_mimi_sweep_cfg4_final.txt
PTTS_PROFILE_PART=4 python profiles/kernel_times.py 2>/dev/null | grep -v WARN > gpurun_out/r02_mimi_kernel_times_final.txt
python - <<'P'
import json
for f in ("r02_bench_final","r02_bench_final2","r02_bench_reference_final","r02_bench_cfg3_final","r02_bench_cfg5_final"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["value"], d.get("ms_per_step"), d.get("e2e", {}).get("value"), d.get("first_chunk"), (d.get("roofline") or {}).get("frac"), d.get("clocks"))
    except Exception as e:
        print(f, "ERR", e)
P
tail -7 gpurun_out/r02_mimi_sweep_cfg4_final.txt
# launch list of the bench command (it exited 0 above without ncu)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r02_launches_bench.csv python bench.py --steps 8 --warmup 4 > gpurun_out/r02_ncu_launches.log 2>&1
echo "ncu launch list exit $?"; wc -l gpurun_out/r02_launches_bench.csv
python profiles/summarize_launches.py gpurun_out/r02_launches_bench.csv > gpurun_out/r02_launches_bench_final.txt 2>&1; head -30 gpurun_out/r02_launches_bench_final.txt
# full capture of the top kernels of the step (eager driver: it exits 0 without ncu first)
python profiles/capture_step.py --eager --frames 8 --warmup 12 > gpurun_out/plain3.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"attn_decode_kernel|flow_head_ss_kernel|attn_window_mma_kernel|gemm_tc_rows_kernel" -s 40 -c 24 -o gpurun_out/r02_final_top_kernels python profiles/capture_step.py --eager --frames 8 --warmup 12 > gpurun_out/ncu3.log 2>&1
echo "ncu full exit $?"; ls -la gpurun_out/r02_final_top_kernels.ncu-rep
