# 64-byte triangle records + two 256-bit loads per step: parity, blocks sweep, soup 1M / 10M
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > gpurun_out/r17_pytest.log; cat gpurun_out/r17_pytest.log
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/r17_bench_$name.json 2> gpurun_out/r17_bench_$name.err; echo "bench $name rc=$?"
  python - <<PY
import json; d=json.load(open("gpurun_out/r17_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), round(d["ms_per_step"],1), "trace ms", round(r["kernel_ms_total"],1), "frac", round(r["frac"],3))
PY
  env "$@" timeout 600 python tools/bench_closest_hit.py --rays 16000000 2>/dev/null | grep '"global"' > gpurun_out/r17_ch_$name.json
  python - <<PY
import json; d=json.loads(open("gpurun_out/r17_ch_$name.json").read()); print("$name soup", round(d["mrays_per_s"],1), "Mrays/s")
PY
}
run b9 RTB_TRACE_BLOCKS=9
run b10 RTB_TRACE_BLOCKS=10
run b9_r8 RTB_TRACE_BLOCKS=9 RTB_TRACE_POLICY=8,1,1
timeout 900 python tools/bench_closest_hit.py --tris 10000000 --half 21.5 --rays 16000000 2>/dev/null | grep '"global"' | tee gpurun_out/r17_ch_10m.json | cut -c1-400
B="python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline"
ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 12 -c 1 -o gpurun_out/r01j_k_wf_trace $B > gpurun_out/ncu_trace.log 2>&1; tail -1 gpurun_out/ncu_trace.log
