#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "stft or spectrogram" 2>&1 | tail -2
timeout 400 python scripts/secondary_bench.py 2> gpurun_out/secondary.err | grep "^{" > gpurun_out/secondary2.jsonl
python - <<'PY'
import json
for l in open('gpurun_out/secondary2.jsonl'):
    d=json.loads(l)
    if 'K4' in d['kernel'] or 'K2' in d['kernel']: print(f"{d['kernel'][:75]:75s} {d['ms']:8.4f} ms {d['GBps']:8.1f} GB/s {d['frac_of_measured_peak']:.3f}")
PY
