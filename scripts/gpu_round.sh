#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -4
timeout 300 python scripts/stft_fm_ab.py > gpurun_out/stft_fm_ab.txt 2>&1; cat gpurun_out/stft_fm_ab.txt
