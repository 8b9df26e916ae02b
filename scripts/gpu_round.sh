#!/bin/bash
mkdir -p gpurun_out
P=$PWD/vector_b200/csrc
for i in 1 2; do
B200VEC_LIB=$P/libb200vec.so timeout 100 python scripts/fused_ab.py LAYOUT=fm 2>&1 | head -1
B200VEC_LIB=$P/libb200vec_tf.so timeout 100 python scripts/fused_ab.py LAYOUT=fm LAYOUT=fm,B200VEC_WF_CHUNK=8 2>&1 | grep -v identical | head -2
B200VEC_LIB=$P/libb200vec.so timeout 100 python scripts/norm_ab.py "" 2>&1 | head -1
B200VEC_LIB=$P/libb200vec_tf.so timeout 100 python scripts/norm_ab.py "" B200VEC_WF_CHUNK=8 2>&1 | grep -v identical | head -2
done
B200VEC_LIB=$P/libb200vec_tf.so timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fused or peer or cfg4 or cfg5 or full_size" 2>&1 | tail -2
