#!/bin/bash
# scratch: one multi-GPU call = bench at N + cfg4 strong scaling at N
N=${1:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 240 $TR bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench rc=$?"; cat gpurun_out/bench_n$N.json | cut -c1-400
timeout 240 $TR scripts/cfg4_scaling.py > gpurun_out/cfg4_n$N.json 2> gpurun_out/cfg4_n$N.err
echo "cfg4 rc=$?"; cat gpurun_out/cfg4_n$N.json
