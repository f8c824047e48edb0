#!/bin/bash
# Round-2 final ncu evidence of the committed kernels (one gpurun call): launch list of the bench step, --set full of
# the four representative conv_tc4 launches (largest launch first: its DRAM bytes are the bench line's `traffic`).
set -x
cd "$(dirname "$0")/.."
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/r2f_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r2f_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/r2f_launch.log 2>&1
python scripts/pconv_probe.py 160,64,120,152,128,3,2,pair,f32 160,128,30,38,128,3,1,pair,f32 160,64,120,152,64,1,1,half,pair 160,32,120,152,32,3,1,pair,pair,32,res 160,32,120,152,64,1,1,pair,half,32 > gpurun_out/r2f_plain2.log 2>&1 || exit 1
i=0
for spec in 160,64,120,152,128,3,2,pair,f32 160,128,30,38,128,3,1,pair,f32 160,64,120,152,64,1,1,half,pair 160,32,120,152,32,3,1,pair,pair,32,res 160,32,120,152,64,1,1,pair,half,32; do
  i=$((i+1))
  timeout -s KILL 400 ncu --set full --clock-control none --import-source on -k regex:conv_tc4 --launch-skip 5 -c 1 -f -o gpurun_out/r2f_tc4_$i \
    python scripts/pconv_probe.py $spec > /dev/null 2>&1
done
ls -la gpurun_out/r2f_*.ncu-rep
