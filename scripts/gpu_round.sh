#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 | tee gpurun_out/pytest_gpu_tail.log
timeout 400 python scripts/secondary_bench.py 2> gpurun_out/secondary.err | grep "^{" > gpurun_out/secondary.jsonl
echo rc=$?
python - <<'PY'
import json
for l in open('gpurun_out/secondary.jsonl'):
    d=json.loads(l); print(f"{d['kernel'][:75]:75s} {d['ms']:8.4f} ms {d['GBps']:8.1f} GB/s {d['frac_of_measured_peak']:.3f}")
PY
