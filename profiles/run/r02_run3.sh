mkdir -p gpurun_out
timeout 120 profiles/probes/l2_ingest_probe2.bin > gpurun_out/r02_l2_ingest_probe2.txt 2>&1; cat gpurun_out/r02_l2_ingest_probe2.txt
PTTS_HEAD_TRACE=1 python -m pytest tests/test_gpu_parity.py -q -s -k "flow_head or ring_wrap or bf16_teacher or bench_shape or cfg3_shape" > gpurun_out/r02_pytest_head.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_pytest_head.log
grep -E "passed|failed|FAILED|Error|assert|rel-L2|SNR|head-ss" gpurun_out/r02_pytest_head.log | tail -40
python profiles/capture_step.py --frames 8 --warmup 4 2>&1 | tail -3
PTTS_HEAD=2 python profiles/capture_step.py --frames 8 --warmup 4 2>&1 | tail -3
