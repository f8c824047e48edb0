#!/bin/bash
# ncu --set full captures of single conv_tc3 launches (one per shape) -> gpurun_out/ncu_<tag>.ncu-rep
i=0
for spec in "$@"; do
  i=$((i+1))
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv_tc3 --launch-skip 2 -c 1 \
    -f -o gpurun_out/ncu_probe_$i python scripts/conv_probe.py --nocheck --iters=1 "$spec" > gpurun_out/ncu_probe_$i.log 2>&1
  echo "$spec -> ncu_probe_$i rc=$?"
done
