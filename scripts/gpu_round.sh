#!/bin/bash
mkdir -p gpurun_out
P=$PWD/vector_b200/csrc
for i in 1 2; do
B200VEC_LIB=$P/libb200vec.so timeout 100 python scripts/fused_ab.py LAYOUT=fm LAYOUT=fm,B200VEC_WF_BULK=0 LAYOUT=fm,B200VEC_WF_BULK=0,B200VEC_WF_PF=1 2>&1 | grep -v identical | head -3
done
