mkdir -p gpurun_out
python profiles/capture_step.py --frames 48 --warmup 12 2>&1 | tail -1
PTTS_PROFILE_PART=3 python profiles/capture_step.py --frames 48 --warmup 12 2>&1 | tail -1
PTTS_PROFILE_PART=4 python profiles/capture_step.py --frames 48 --warmup 12 2>&1 | tail -1
PTTS_PROFILE_PART=3 PTTS_TIMELINE=1 python profiles/capture_step.py --frames 24 --warmup 12 > gpurun_out/tl_flow_only.txt 2>&1
python profiles/timeline_summary.py gpurun_out/tl_flow_only.txt
PTTS_TIMELINE=1 python profiles/capture_step.py --frames 24 --warmup 12 > gpurun_out/tl_full.txt 2>&1
python profiles/timeline_summary.py gpurun_out/tl_full.txt
