# L2 persistence window over the per-hit records (nodes, triangles, shading records, tangents): A/B
mkdir -p gpurun_out
python - <<'PY'
import torch, ctypes
rt = ctypes.CDLL("libcudart.so.12") if False else None
p = torch.cuda.get_device_properties(0)
print("L2", p.L2_cache_size, "persisting max", getattr(p, "persisting_l2_cache_max_size", None), "window max", getattr(p, "access_policy_max_window_size", None))
PY
python -m pytest tests/test_gpu_render.py -m gpu -x -q 2>&1 | tail -2
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline $SIZE > gpurun_out/r30_bench_$name.json 2> gpurun_out/r30_bench_$name.err
  python - <<PY
import json; d=json.load(open("gpurun_out/r30_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; trace", round(r["kernel_ms_total"],1), "other", round(r["other_kernels_ms_total"],1))
PY
}
SIZE=""; run full_on; run full_off RTB_L2_PERSIST=0
SIZE="--width 960 --height 272"; run eighth_on; run eighth_off RTB_L2_PERSIST=0
B="python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:k_wf_shade -s 14 -c 1 $B 2>&1 | grep -E "duration|dram__bytes|hit_rate" | head -4
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:k_wf_trace -s 14 -c 1 $B 2>&1 | grep -E "duration|dram__bytes|hit_rate" | head -4
