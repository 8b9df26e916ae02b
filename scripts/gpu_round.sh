#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/norm_ab.py B200VEC_WF_ZW=1 B200VEC_WF_ZW=-1 B200VEC_WF_ZW=0 B200VEC_WF_ZW=-1 B200VEC_WF_ZW=2 B200VEC_WF_ZW=1 2>&1 | grep -v identical | tail -12
