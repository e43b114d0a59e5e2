# L2 prefetch of the shading data (texture taps, tangents, normal transform) in the shade step
mkdir -p gpurun_out
python -m pytest tests/test_gpu_render.py -m gpu -x -q 2>&1 | tail -2
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline $SIZE > gpurun_out/r29_bench_$name.json 2> gpurun_out/r29_bench_$name.err
  python - <<PY
import json; d=json.load(open("gpurun_out/r29_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; trace", round(r["kernel_ms_total"],1), "other", round(r["other_kernels_ms_total"],1))
PY
}
SIZE=""; run full
run full_s8 RTB_SHADE_BLOCKS=8
run full_s5 RTB_SHADE_BLOCKS=5
SIZE="--width 960 --height 272"; run eighth
