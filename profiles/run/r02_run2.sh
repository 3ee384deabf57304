mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_pytest_gpu.log
grep -E "passed|failed|FAILED|Error|assert" gpurun_out/r02_pytest_gpu.log | tail -30
timeout 120 profiles/probes/l2_ingest_probe.bin > gpurun_out/r02_l2_ingest_probe.txt 2>&1; cat gpurun_out/r02_l2_ingest_probe.txt
