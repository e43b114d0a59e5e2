# form 3 after the register diet (uv in local memory, 32-bit ray counter, 64-bit stack entries): parity, blocks sweep, ncu of the trace kernel
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > gpurun_out/r16_pytest.log; cat gpurun_out/r16_pytest.log
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/r16_bench_$name.json 2> gpurun_out/r16_bench_$name.err; echo "bench $name rc=$?"
  python - <<PY
import json; d=json.load(open("gpurun_out/r16_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), round(d["ms_per_step"],1), "trace ms", round(r["kernel_ms_total"],1), "frac", round(r["frac"],3))
PY
  env "$@" timeout 600 python tools/bench_closest_hit.py --rays 16000000 2>/dev/null | grep '"global"' > gpurun_out/r16_ch_$name.json
  python - <<PY
import json; d=json.loads(open("gpurun_out/r16_ch_$name.json").read()); print("$name soup", round(d["mrays_per_s"],1), "Mrays/s")
PY
}
run m3_b9 RTB_TRACE_BLOCKS=9
run m3_b10 RTB_TRACE_BLOCKS=10
run m3_b10_r2 RTB_TRACE_BLOCKS=10 RTB_TRACE_POLICY=2,1,1
run m3_b10_r8 RTB_TRACE_BLOCKS=10 RTB_TRACE_POLICY=8,1,1
B="env RTB_TRACE_BLOCKS=10 python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 12 -c 1 -o gpurun_out/r01i_k_wf_trace $B > gpurun_out/ncu_trace.log 2>&1; tail -1 gpurun_out/ncu_trace.log
