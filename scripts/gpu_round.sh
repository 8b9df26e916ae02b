#!/bin/bash
mkdir -p gpurun_out
for i in 1 2; do
timeout 300 python scripts/norm_ab.py "" B200VEC_WF_ZW=0 B200VEC_WF_ZW=1 B200VEC_WF_ZW=2 "" 2>&1 | tail -9
done
B200VEC_WF_ZW=1 timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fused or peer or cfg4 or cfg5" 2>&1 | tail -2
