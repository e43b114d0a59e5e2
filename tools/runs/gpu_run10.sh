mkdir -p gpurun_out
nvidia-smi -L
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tools/check_multigpu.py > gpurun_out/check_multigpu.log 2>&1; echo "check rc=$?"; tail -2 gpurun_out/check_multigpu.log
timeout 900 $TR bench.py --gpus 2 --spp 128 --steps 2 --warmup 3 > gpurun_out/bench_n2_spp128.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"; cat gpurun_out/bench_n2_spp128.json | cut -c1-1800; tail -3 gpurun_out/bench_n2.err
timeout 900 python bench.py --gpus 1 --spp 128 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1_spp128.json 2> gpurun_out/bench_n1.err; echo "bench n1 rc=$?"; cat gpurun_out/bench_n1_spp128.json | cut -c1-700
python scenes/gen_scenes.py sponza > /dev/null 2>&1
RT_STATS=1 RT_GPUS=2 ./run.sh scenes/_gen/sponza/sponza.gltf 960 540 32 gpurun_out/cli_n2.ppm > gpurun_out/cli_n2.log 2>&1; echo "cli n2 rc=$?"; tail -2 gpurun_out/cli_n2.log
RT_STATS=1 RT_GPUS=1 ./run.sh scenes/_gen/sponza/sponza.gltf 960 540 32 gpurun_out/cli_n1.ppm > gpurun_out/cli_n1.log 2>&1; echo "cli n1 rc=$?"; tail -2 gpurun_out/cli_n1.log
cmp gpurun_out/cli_n1.ppm gpurun_out/cli_n2.ppm && echo "CLI images identical"; rm -f gpurun_out/cli_n1.ppm
