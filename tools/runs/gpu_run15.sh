# MODE 3 (one pop / decode site, box_test_z) vs MODE 1: parity, then closest-hit microbench + 64-spp frame per (mode, blocks)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > gpurun_out/r15_pytest.log; cat gpurun_out/r15_pytest.log
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/r15_bench_$name.json 2> gpurun_out/r15_bench_$name.err; echo "bench $name rc=$?"
  python - <<PY
import json; d=json.load(open("gpurun_out/r15_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), round(d["ms_per_step"],1), "trace ms", round(r["kernel_ms_total"],1), "frac", round(r["frac"],3))
PY
  env "$@" timeout 600 python tools/bench_closest_hit.py --rays 16000000 2>/dev/null | grep '"global"' > gpurun_out/r15_ch_$name.json
  python - <<PY
import json; d=json.loads(open("gpurun_out/r15_ch_$name.json").read()); print("$name soup", round(d["mrays_per_s"],1), "Mrays/s")
PY
}
run m1_b10 RTB_TRACE_MODE=1 RTB_TRACE_BLOCKS=10
run m3_b10 RTB_TRACE_MODE=3 RTB_TRACE_BLOCKS=10
run m3_b9 RTB_TRACE_MODE=3 RTB_TRACE_BLOCKS=9
run m3_b12 RTB_TRACE_MODE=3 RTB_TRACE_BLOCKS=12
