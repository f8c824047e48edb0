#!/bin/bash
# ncu --set full (with source) of single conv_tc4 launches, one per pconv_probe spec -> gpurun_out/ncu_pconv_<i>.ncu-rep
# (each spec is first run WITHOUT ncu and must exit 0)
cd "$(dirname "$0")/.."
i=0
for spec in "$@"; do
  i=$((i+1))
  timeout -s KILL 120 python scripts/pconv_probe.py "$spec" > gpurun_out/ncu_pconv_$i.plain.log 2>&1 || { echo "$spec: plain run failed"; continue; }
  timeout -s KILL 400 ncu --set full --clock-control none --import-source on -k regex:conv_tc4 --launch-skip 3 -c 1 \
    -f -o gpurun_out/ncu_pconv_$i python scripts/pconv_probe.py "$spec" > gpurun_out/ncu_pconv_$i.log 2>&1
  echo "$spec -> ncu_pconv_$i rc=$?"
done
