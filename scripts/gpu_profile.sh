#!/bin/bash
# ncu evidence for the bench step: launch list + full capture of the hot kernel (each only after the
# same command has exited 0 without ncu).  Outputs land in gpurun_out/.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-cfg4 --no-cfg5 --no-save --load-steps 20"
KREGEX="${KREGEX:-wf1024}"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$KREGEX" -s 3 -c 1 -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | head; tail -2 gpurun_out/ncu_full.log
