#!/bin/bash
# 8-GPU runs of the final code: cfg5 (2048 streams, generate_audio_batch) and the default cfg2 bench, one rank per GPU.
set -x
mkdir -p gpurun_out
P=29517
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $P bench.py --gpus 8 --workload cfg5 > gpurun_out/r02_bench_cfg5_8gpu.json 2> gpurun_out/r02_bench_cfg5_8gpu.err
echo "cfg5 exit $?"; tail -c 1500 gpurun_out/r02_bench_cfg5_8gpu.json
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((P+1)) bench.py --gpus 8 > gpurun_out/r02_bench_8gpu.json 2> gpurun_out/r02_bench_8gpu.err
echo "cfg2 exit $?"; tail -c 1500 gpurun_out/r02_bench_8gpu.json
