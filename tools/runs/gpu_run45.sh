# default build (batch kernel with the 8-entry ring, frame kernel without): parity; then the second pop entry read only when the first is culled
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/r45_bench_$name.json 2> gpurun_out/r45_bench_$name.err
  python - <<PY
import json; d=json.load(open("gpurun_out/r45_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; trace", round(r["kernel_ms_total"],1), "other", round(r["other_kernels_ms_total"],1))
PY
  env "$@" python tools/bench_closest_hit.py --rays 16000000 2>/dev/null | grep '"global"' | cut -c1-150
}
run default A=1
cp raytracing-hw_b200/variants/librtb_lazy.so raytracing-hw_b200/librtb.so
run lazy A=1
run lazy_again A=1
