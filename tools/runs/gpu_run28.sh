# evict-first (ld/st.cs) on the path-state arrays: frame rate full / quarter / eighth canvas, shade kernel DRAM traffic
mkdir -p gpurun_out
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --spp 64 --steps 3 --warmup 2 --no-cpu-baseline $SIZE > gpurun_out/r28_bench_$name.json 2> gpurun_out/r28_bench_$name.err
  python - <<PY
import json; d=json.load(open("gpurun_out/r28_bench_$name.json")); r=d["roofline"]; print("$name", round(d["value"],1), "Mpix*spp/s", round(d["ms_per_step"],1), "ms; trace", round(r["kernel_ms_total"],1), "other", round(r["other_kernels_ms_total"],1))
PY
}
SIZE=""; run full
SIZE="--width 960 --height 544"; run quarter
SIZE="--width 960 --height 272"; run eighth
B="python bench.py --spp 8 --steps 1 --warmup 1 --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:k_wf_shade -s 14 -c 3 $B 2>&1 | grep -E "k_wf_shade|duration|dram__bytes|hit_rate" | head -16
