# slot groups on the small scene of cfg2 (s7_4 stand-in, 53 BVH nodes): the frame is shade-heavy there
mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 300 python bench.py --workload s7_4 --steps 3 --warmup 2 --no-cpu-baseline $SIZE > gpurun_out/r48_bench_$name.json 2> gpurun_out/r48_bench_$name.err
  python - <<PY
import json; d=json.load(open("gpurun_out/r48_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; trace", round(r["kernel_ms_total"],1), "other", round(r["other_kernels_ms_total"],1), "launches", d["gpu_launches"])
PY
}
SIZE=""
for g in 1 2 3 4 6; do run cfg2_g$g RTB_WF_GROUPS=$g; done
SIZE="--width 1920 --height 1080 --spp 64"
for g in 1 2 4; do run hd_g$g RTB_WF_GROUPS=$g; done
SIZE="--width 640 --height 360 --spp 256"
for g in 1 2 4; do run small_g$g RTB_WF_GROUPS=$g; done
