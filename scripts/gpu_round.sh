#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -4
python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/bench_one.json 2> gpurun_out/bench_one.err; tail -2 gpurun_out/bench_one.err
python bench.py --steps 20 --warmup 3 --no-cpu --two-launch > gpurun_out/bench_two.json 2>&1
python - <<'PY'
import json
for f in ("bench_one","bench_two"):
    d=json.load(open(f"gpurun_out/{f}.json"))
    print(f, d["value"], d["ms_per_step"], {k:v["ms"] for k,v in d["kernels"].items()}, d["roofline"]["frac"], d["e2e"]["value"], d["gpu_launches"])
PY
