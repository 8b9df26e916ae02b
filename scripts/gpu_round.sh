#!/bin/bash
mkdir -p gpurun_out
for i in 1 2; do
for lib in prev ""; do
  L=$PWD/vector_b200/csrc/libb200vec${lib:+_$lib}.so
  echo "== $L"
  B200VEC_LIB=$L timeout 120 python scripts/fused_ab.py LAYOUT=fm 2>&1 | grep -v identical | head -1
  B200VEC_LIB=$L timeout 120 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/bx.json 2> gpurun_out/bx.err
  python - <<PY
import json
d=json.load(open("gpurun_out/bx.json")); print("bench", round(d["value"],1), round(d["ms_per_step"]*1e3,1), {k:round(v["ms"]*1e3,1) for k,v in d["kernels"].items()})
PY
done; done
