mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
for pol in 8,1,1 8,2,5 4,2,5 16,2,5 8,1,2 8,1,4; do
  echo "policy $pol"; RTB_TRACE_POLICY=$pol timeout 300 python tools/bench_closest_hit.py --reps 2 2>&1 | grep global | tee -a gpurun_out/ch_policy.jsonl
done
for blk in 6 10 12; do
  echo "blocks $blk"; RTB_TRACE_BLOCKS=$blk timeout 300 python tools/bench_closest_hit.py --reps 2 2>&1 | grep global | tee -a gpurun_out/ch_blocks.jsonl
done
timeout 600 python bench.py --spp 64 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_wf_spp64.json 2> gpurun_out/bench_wf_spp64.err; echo "bench wf rc=$?"; cat gpurun_out/bench_wf_spp64.json; tail -3 gpurun_out/bench_wf_spp64.err
RTB_TRACE_POLICY=8,2,5 timeout 600 python bench.py --spp 64 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_wf_spp64_p825.json 2>&1; cat gpurun_out/bench_wf_spp64_p825.json
timeout 600 python bench.py --spp 64 --steps 2 --warmup 1 --no-cpu-baseline --megakernel > gpurun_out/bench_mk_spp64.json 2> gpurun_out/bench_mk_spp64.err; echo "bench mk rc=$?"; cat gpurun_out/bench_mk_spp64.json
