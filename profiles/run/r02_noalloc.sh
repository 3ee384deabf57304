mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/r02_pytest_noalloc_alone.log 2>&1; echo "alone exit $?" | tee -a gpurun_out/r02_pytest_noalloc_alone.log
tail -3 gpurun_out/r02_pytest_noalloc_alone.log
python -m pytest tests -x -q -m gpu > gpurun_out/r02_pytest_gpu.log 2>&1; echo "suite exit $?" | tee -a gpurun_out/r02_pytest_gpu.log
grep -E "passed|failed|FAILED|Error" gpurun_out/r02_pytest_gpu.log | tail -8
python __graft_entry__.py --smoke 2>&1 | grep -a smoke
python profiles/capture_step.py --eager --frames 1 --warmup 2 > gpurun_out/plain_capture.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 78 -c 26 -o gpurun_out/r02_skinny_gemm python profiles/capture_step.py --eager --frames 1 --warmup 2 > gpurun_out/ncu_capture.log 2>&1
ncu -i gpurun_out/r02_skinny_gemm.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,launch__grid_size,launch__cluster_dim_z,dram__bytes_read.sum,dram__bytes_write.sum,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,smsp__cycles_active.avg,sm__throughput.avg.pct_of_peak_sustained_elapsed,launch__grid_dim_x,launch__grid_dim_y,launch__grid_dim_z > gpurun_out/r02_skinny_gemm_ncu_summary.csv 2>&1
wc -l gpurun_out/r02_skinny_gemm_ncu_summary.csv
