#!/bin/bash
# scratch: bench lines of the final library at N = 8 and 4 on one 8-GPU box + the peer-exchange check
mkdir -p gpurun_out
for N in 8 4; do
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N"
timeout 200 $TR bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_v9_n$N.json 2> gpurun_out/bench_v9_n$N.err
echo "bench N=$N rc=$?"; cut -c1-260 gpurun_out/bench_v9_n$N.json
done
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533"
timeout 200 $TR scripts/peer_check.py > gpurun_out/peer_check_v9_n8.txt 2>&1; echo "peer rc=$?"; tail -3 gpurun_out/peer_check_v9_n8.txt
