#!/bin/bash
timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv_direct --launch-skip 1 -c 1 \
  -f -o gpurun_out/ncu_direct python scripts/conv_probe.py direct:32,5,240,304,64,3,2 > gpurun_out/ncu_direct.log 2>&1
echo rc=$?
