mkdir -p gpurun_out
export PTTS_PROFILE_PART=0
python profiles/capture_step.py --eager --frames 8 --warmup 12 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"seanet_resblock_fused|attn_prefill_mma" -c 6 -o gpurun_out/r02_resblock_prefill python profiles/capture_step.py --eager --frames 8 --warmup 12 > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out/r02_resblock_prefill.ncu-rep
