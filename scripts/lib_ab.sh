#!/bin/bash
# Same-box A/B of whole libraries on the one-launch step: scripts/lib_ab.sh libA.so libB.so ...  (run twice, interleaved)
for rep in 1 2; do
  for lib in "$@"; do
    B200VEC_LIB=$lib python scripts/norm_ab.py "" 2>&1 | grep "us/step"
  done
done
