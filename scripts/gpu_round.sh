#!/bin/bash
mkdir -p gpurun_out
python scripts/fused_ab.py LAYOUT=fm LAYOUT=fm,B200VEC_WF_TAIL=0 LAYOUT=fm,B200VEC_WF_CHUNK=8 LAYOUT=fm,B200VEC_WF_CHUNK=2 LAYOUT=fm,B200VEC_WF_CHUNK=2,B200VEC_WF_TAIL=0 LAYOUT=fm,B200VEC_WF_BULK=0 LAYOUT=fm,B200VEC_WF_BULK=0,B200VEC_WF_CHUNK=2 > gpurun_out/wf_ab.txt 2>&1
timeout 600 python -m pytest tests -m gpu -q -x -k "fused" 2>&1 | tail -3
cat gpurun_out/wf_ab.txt
