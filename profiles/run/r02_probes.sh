mkdir -p gpurun_out
timeout 120 profiles/probes/mma_sync_probe.bin > gpurun_out/r02_mma_sync_probe.txt 2>&1; cat gpurun_out/r02_mma_sync_probe.txt
timeout 120 profiles/probes/l2_ingest_probe2.bin > gpurun_out/r02_l2_ingest_probe2b.txt 2>&1; grep -E "lanes|warps 8|64x256" gpurun_out/r02_l2_ingest_probe2b.txt
