# 2- and 4-rank lines of the default workload (config 2) for profiles/r2_configs.md; run under gpurun --gpus 4
set -x
mkdir -p gpurun_out/cfg
for n in 2 4; do
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n"
  timeout -s KILL 300 $TR bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/cfg/r2_bench_config2_${n}gpu.json 2> gpurun_out/cfg/c2_$n.err
  tail -n 3 gpurun_out/cfg/c2_$n.err
done
cat gpurun_out/cfg/r2_bench_config2_[24]gpu.json | cut -c1-300
