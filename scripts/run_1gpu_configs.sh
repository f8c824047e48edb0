# Every one-GPU line of profiles/r2_configs.md (run under gpurun, one GPU): bench.py output -> gpurun_out/cfg/
set -x
mkdir -p gpurun_out/cfg
B="timeout -s KILL 400 python bench.py"
$B > gpurun_out/cfg/r2_bench_config2.json 2> gpurun_out/cfg/c2.err
$B --impl reference --steps 3 --warmup 1 > gpurun_out/cfg/r2_bench_config2_reference_arm.json 2> gpurun_out/cfg/c2r.err
$B --config 1 > gpurun_out/cfg/r2_bench_config1.json 2> gpurun_out/cfg/c1.err
$B --config 1 --batch 64 --steps 20 --no-cpu-baseline > gpurun_out/cfg/r2_bench_config1_b64.json 2> gpurun_out/cfg/c1b.err
$B --config 3 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/cfg/r2_bench_config3_1gpu.json 2> gpurun_out/cfg/c3.err
$B --config 5 --batch 64 --T 16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/cfg/r2_bench_config5_b64_t16.json 2> gpurun_out/cfg/c5a.err
$B --config 5 --batch 64 --T 32 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/cfg/r2_bench_config5_b64_t32.json 2> gpurun_out/cfg/c5b.err
$B --config 5 --batch 64 --T 64 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/cfg/r2_bench_config5_b64_t64.json 2> gpurun_out/cfg/c5c.err
$B --config 5 --batch 256 --T 32 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/cfg/r2_bench_config5_b256_t32.json 2> gpurun_out/cfg/c5d.err
$B --config 5 --batch 1024 --T 16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/cfg/r2_bench_config5_b1024_t16.json 2> gpurun_out/cfg/c5e.err
timeout -s KILL 300 python scripts/train_step_probe.py --steps 5 > gpurun_out/cfg/r2_train_1gpu.json 2> gpurun_out/cfg/tr.err
tail -n 2 gpurun_out/cfg/*.err | cut -c1-200
for f in gpurun_out/cfg/*.json; do echo $f; tail -n 1 $f | cut -c1-200; done
