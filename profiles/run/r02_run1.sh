mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_pytest_gpu.log
python __graft_entry__.py --smoke > gpurun_out/r02_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r02_smoke.log
python bench.py > gpurun_out/r02_bench0.json 2> gpurun_out/r02_bench0.err
python profiles/capture_step.py --eager --frames 2 --warmup 2 > gpurun_out/plain_capture.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 60 -c 30 -o gpurun_out/r02_skinny_gemm python profiles/capture_step.py --eager --frames 2 --warmup 2 > gpurun_out/ncu_capture.log 2>&1
tail -3 gpurun_out/r02_pytest_gpu.log; cat gpurun_out/r02_smoke.log | tail -4; cat gpurun_out/r02_bench0.json | head -c 1500
