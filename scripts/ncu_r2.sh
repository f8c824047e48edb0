#!/bin/bash
# Round-2 ncu evidence (one gpurun call): launch list of the bench step, --set full of the dominant conv_tc4 launch
# shapes and of the HBM-bound kernels.  Summaries are extracted here with `ncu -i ... --page raw --csv`.
set -x
cd "$(dirname "$0")/.."
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/r2_ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/r2_ncu_launch.log 2>&1
python scripts/pconv_probe.py 160,64,120,152,128,3,2,pair,f32 160,128,30,38,128,3,1,pair,f32 160,64,120,152,64,1,1,half,pair 160,32,120,152,32,3,1,pair,pair,32,res > gpurun_out/r2_ncu_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:conv_tc4 -s 12 -c 1 -o gpurun_out/r2_tc4_s2 \
    python scripts/pconv_probe.py 160,64,120,152,128,3,2,pair,f32 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:conv_tc4 -s 12 -c 1 -o gpurun_out/r2_tc4_128 \
    python scripts/pconv_probe.py 160,128,30,38,128,3,1,pair,f32 > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:conv_tc4 -s 12 -c 1 -o gpurun_out/r2_tc4_1x1 \
    python scripts/pconv_probe.py 160,64,120,152,64,1,1,half,pair > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:conv_tc4 -s 7 -c 1 -o gpurun_out/r2_tc4_lif \
    python scripts/pconv_probe.py 160,32,120,152,32,3,1,pair,pair,32,res > /dev/null 2>&1
python scripts/scan_probe.py 32,5,64,60,76,pair,pair > gpurun_out/r2_ncu_plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:lif_fmt -s 12 -c 1 -o gpurun_out/r2_lif_fmt \
    python scripts/scan_probe.py 32,5,64,60,76,pair,pair > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k "regex:bin_bitmap|conv3_lif_pool2|im2col3|li_|neuron_fwd" -s 40 -c 12 -o gpurun_out/r2_hbm_kernels \
    python bench.py --config 1 --batch 64 --steps 1 --warmup 3 --no-cpu-baseline --no-graph > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k "regex:bin_bitmap|im2col3|neuron_fwd" -s 12 -c 6 -o gpurun_out/r2_hbm_kernels2 \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > /dev/null 2>&1
ls -la gpurun_out/*.ncu-rep
