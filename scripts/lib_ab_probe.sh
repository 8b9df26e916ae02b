#!/bin/bash
# Same-box A/B of whole libraries on scripts/norm_build_probe.py: scripts/lib_ab_probe.sh libA.so libB.so ...
for lib in "$@"; do
  echo "== $lib"; B200VEC_LIB=$lib python scripts/norm_build_probe.py 2>&1 | grep " us$" | tail -6
done
