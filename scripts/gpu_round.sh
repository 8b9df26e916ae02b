#!/bin/bash
mkdir -p gpurun_out
P=$PWD/vector_b200/csrc
for i in 1 2; do
B200VEC_LIB=$P/libb200vec.so timeout 100 python scripts/norm_ab.py "" 2>&1 | head -1
B200VEC_LIB=$P/libb200vec_pd.so timeout 100 python scripts/norm_ab.py "" 2>&1 | head -1
done
B200VEC_LIB=$P/libb200vec_pd.so timeout 200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fused" 2>&1 | tail -2
