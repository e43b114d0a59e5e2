mkdir -p gpurun_out
nvidia-smi -L | wc -l
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
timeout 600 $TR tools/check_multigpu.py > gpurun_out/check_multigpu_n$N.log 2>&1; echo "check rc=$?"; tail -1 gpurun_out/check_multigpu_n$N.log
timeout 900 $TR bench.py --gpus $N --spp 256 --steps 2 --warmup 3 > gpurun_out/bench_n${N}_spp256.json 2> gpurun_out/bench_n$N.err; echo "bench n$N rc=$?"; cat gpurun_out/bench_n${N}_spp256.json | cut -c1-1500; tail -2 gpurun_out/bench_n$N.err
timeout 300 $TR bench.py --impl reference --gpus $N --steps 1 --warmup 1 --cpu-sample 480x270x8 > gpurun_out/bench_ref_n$N.json 2>/dev/null; cat gpurun_out/bench_ref_n$N.json | cut -c1-300
