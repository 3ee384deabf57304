mkdir -p gpurun_out
export PTTS_PROFILE_PART=4
python profiles/capture_step.py --eager --frames 8 --warmup 12 > gpurun_out/mimi_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__grid_size,launch__registers_per_thread,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed --clock-control none -c 1200 --csv --log-file gpurun_out/r02_mimi_pass_ncu.csv python profiles/capture_step.py --eager --frames 8 --warmup 12 > gpurun_out/mimi_ncu.log 2>&1
tail -2 gpurun_out/mimi_plain.log; wc -l gpurun_out/r02_mimi_pass_ncu.csv
