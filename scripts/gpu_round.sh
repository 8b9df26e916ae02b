#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 240 $TR bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench rc=$?"; cut -c1-330 gpurun_out/bench_n$N.json
timeout 240 $TR scripts/peer_check.py > gpurun_out/peer_check_n$N.txt 2>&1
echo "peer rc=$?"; tail -4 gpurun_out/peer_check_n$N.txt
