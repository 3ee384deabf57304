mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -q -s -k "flow_head or bf16_teacher" > gpurun_out/r02_pytest_head.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_pytest_head.log
grep -E "passed|failed|FAILED|Error|assert " gpurun_out/r02_pytest_head.log | tail -10
PTTS_HEAD_TRACE=1 python profiles/capture_step.py --eager --frames 8 --warmup 4 2>&1 | grep -a -E "trace|per frame" | tail -4
python profiles/capture_step.py --frames 24 --warmup 12 2>&1 | tail -1
PTTS_HEAD=2 python profiles/capture_step.py --frames 24 --warmup 12 2>&1 | tail -1
PTTS_PROFILE_PART=3 python profiles/capture_step.py --frames 24 --warmup 12 2>&1 | tail -1
PTTS_PROFILE_PART=3 PTTS_HEAD=2 python profiles/capture_step.py --frames 24 --warmup 12 2>&1 | tail -1
