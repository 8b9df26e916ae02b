#!/bin/bash
# One GPU-box visit of experiments: store-pattern floor, fused A/B, phase accounting, ncu full of the fused kernel.
mkdir -p gpurun_out
scripts/ubench/store_patterns > gpurun_out/store_patterns.txt 2>&1
python scripts/fused_ab.py "" B200VEC_FUSED_W=8 B200VEC_FUSED_V=16 B200VEC_FUSED_ZTMA=1 > gpurun_out/fused_ab.txt 2>&1
python scripts/pipe_timing.py > gpurun_out/pipe_timing_w16.txt 2>&1
B200VEC_FUSED_W=8 python scripts/pipe_timing.py > gpurun_out/pipe_timing_w8.txt 2>&1
CMD="python bench.py --steps 2 --warmup 3 --no-cpu"
ncu --set full --clock-control none --import-source on -k regex:tile1024 -s 3 -c 1 -o gpurun_out/prof_fused $CMD > gpurun_out/ncu_fused.log 2>&1
tail -4 gpurun_out/fused_ab.txt; tail -12 gpurun_out/pipe_timing_w16.txt
