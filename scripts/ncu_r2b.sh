#!/bin/bash
# Round-2 ncu evidence, part 2: config-1 kernels (--set full) and the whole-step DRAM traffic of configs 1 and 2.
set -x
cd "$(dirname "$0")/.."
python bench.py --config 1 --batch 64 --steps 1 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/r2_ncu_plain4.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:conv3_lif_pool2|conv_tc4" -s 9 -c 3 -o gpurun_out/r2_config1_kernels \
    python bench.py --config 1 --batch 64 --steps 1 --warmup 3 --no-cpu-baseline --no-graph > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_config1_dram.csv \
    python bench.py --config 1 --batch 64 --steps 1 --warmup 3 --no-cpu-baseline --no-graph > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r2_config2_dram.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-graph > /dev/null 2>&1
ls -la gpurun_out/r2_config*
