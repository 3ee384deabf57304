#!/bin/bash
# cfg2 at N ranks of one box (one process per GPU, gloo control plane, no data-path collective)
N=$1
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N > gpurun_out/r02_bench_${N}gpu.json 2> gpurun_out/r02_bench_${N}gpu.err
echo "exit $?"; python -c "
import json
d=json.loads(open('gpurun_out/r02_bench_${N}gpu.json').read().strip().splitlines()[-1]); print($N, d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks'])"
