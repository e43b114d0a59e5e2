# 8-GPU box: N-GPU image == 1-GPU image over NCCL, then the bench at N = 8, 4, 2 (256 spp) and N = 8 at the full 1024 spp
mkdir -p gpurun_out
nvidia-smi -L | head -8
TR() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 "$@"; }
timeout 600 bash -c "$(declare -f TR); TR 8 tools/check_multigpu.py" > gpurun_out/r01k_check_multigpu_n8.log 2>&1; echo "check rc=$?"; tail -2 gpurun_out/r01k_check_multigpu_n8.log
for n in 8 4 2; do
  timeout 900 bash -c "$(declare -f TR); TR $n bench.py --gpus $n --spp 256 --steps 2 --warmup 3 --no-cpu-baseline" > gpurun_out/r01k_bench_n${n}_spp256.json 2> gpurun_out/r26_bench_n$n.err; echo "bench n$n rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r01k_bench_n${n}_spp256.json')); print($n, round(d['value'],1), 'Mpix*spp/s', round(d['ms_per_step'],1), 'ms', round(d['mrays_per_s'],1), 'Mrays/s e2e', round(d['e2e']['value'],1))"
done
timeout 900 python bench.py --gpus 1 --spp 256 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r01k_bench_n1_spp256.json 2> gpurun_out/r26_bench_n1.err
python -c "
import json; d=json.load(open('gpurun_out/r01k_bench_n1_spp256.json')); print(1, round(d['value'],1), 'Mpix*spp/s', round(d['ms_per_step'],1), 'ms')"
timeout 900 bash -c "$(declare -f TR); TR 8 bench.py --gpus 8 --steps 3 --warmup 3 --no-cpu-baseline" > gpurun_out/r01k_bench_n8_full.json 2> gpurun_out/r26_bench_n8_full.err; echo "bench n8 full rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r01k_bench_n8_full.json')); print('8 full', round(d['value'],1), 'Mpix*spp/s', round(d['ms_per_step'],1), 'ms', round(d['mrays_per_s'],1), 'Mrays/s')"
