# leaf state inside the link (no tri_i / tri_end registers, no decode): parity, blocks 9 / 10
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline $SIZE > gpurun_out/r40_bench_$name.json 2> gpurun_out/r40_bench_$name.err
  python - <<PY
import json; d=json.load(open("gpurun_out/r40_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; trace", round(r["kernel_ms_total"],1), "other", round(r["other_kernels_ms_total"],1))
PY
}
SIZE=""; run full_b9 RTB_TRACE_BLOCKS=9; run full_b10 RTB_TRACE_BLOCKS=10
SIZE="--width 960 --height 272"; run eighth_b9 RTB_TRACE_BLOCKS=9; run eighth_b10 RTB_TRACE_BLOCKS=10
python tools/bench_closest_hit.py --rays 16000000 2>/dev/null | grep '"global"' | cut -c1-200
