set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout -s KILL 300 $TR bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/cfg/r2_bench_config2_8gpu.json 2> gpurun_out/cfg/c2_8.err
timeout -s KILL 300 $TR bench.py --gpus 8 --config 3 --steps 5 --warmup 3 > gpurun_out/cfg/r2_bench_config3_8gpu.json 2> gpurun_out/cfg/c3_8.err
timeout -s KILL 300 $TR bench.py --gpus 8 --config 5 --batch 1024 --T 32 --steps 2 --warmup 3 > gpurun_out/cfg/r2_bench_config5_b1024_t32_8gpu.json 2> gpurun_out/cfg/c5_8.err
tail -n 3 gpurun_out/cfg/c2_8.err gpurun_out/cfg/c3_8.err gpurun_out/cfg/c5_8.err
cat gpurun_out/cfg/*8gpu.json | cut -c1-300
