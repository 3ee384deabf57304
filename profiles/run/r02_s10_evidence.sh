mkdir -p gpurun_out
python profiles/mimi_sweep.py > gpurun_out/r02_mimi_sweep_cfg4.txt 2>&1; cat gpurun_out/r02_mimi_sweep_cfg4.txt | tail -7
PTTS_PROFILE_PART=4 python profiles/kernel_times.py --frames 32 > gpurun_out/r02_mimi_kernel_times.txt 2>&1; head -24 gpurun_out/r02_mimi_kernel_times.txt
export PTTS_PROFILE_PART=4
python profiles/capture_step.py --eager --frames 8 --warmup 12 > gpurun_out/mimi_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:seanet_stage2_fused -s 2 -c 2 -o gpurun_out/r02_fused_stage2 python profiles/capture_step.py --eager --frames 8 --warmup 12 > gpurun_out/mimi_ncu.log 2>&1
ls -la gpurun_out/*.ncu-rep
