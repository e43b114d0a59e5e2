# 8-GPU box, final code: image identity, then the default bench (1024 spp) at N = 8, 4, 2
mkdir -p gpurun_out
TR() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 "$@"; }
timeout 600 bash -c "$(declare -f TR); TR 8 tools/check_multigpu.py" > gpurun_out/r01l_check_multigpu_n8.log 2>&1; echo "check rc=$?"; tail -1 gpurun_out/r01l_check_multigpu_n8.log
for n in 8 4 2; do
  timeout 900 bash -c "$(declare -f TR); TR $n bench.py --gpus $n --steps 2 --warmup 3 --no-cpu-baseline" > gpurun_out/r01l_bench_n${n}_full.json 2> gpurun_out/r39_bench_n$n.err; echo "bench n$n rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/r01l_bench_n${n}_full.json')); print($n, round(d['value'],1), 'Mpix*spp/s', round(d['ms_per_step'],1), 'ms', round(d['mrays_per_s'],1), 'Mrays/s e2e', round(d['e2e']['value'],1))"
done
