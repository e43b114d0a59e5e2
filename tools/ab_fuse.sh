for i in 1 2; do
for f in 1 0; do
echo "--- RTB_FUSE_TEXTURES=$f"
RTB_FUSE_TEXTURES=$f python bench.py --steps 3 --warmup 3 --spp 64 --no-cpu-baseline --no-e2e 2>/dev/null | python -c 'import sys,json; d=json.loads(sys.stdin.read()); r=d["roofline"]; print(json.dumps({"value": d["value"], "trace_ms": r["kernel_ms_total"], "other_ms": r["other_kernels_ms_total"], "kernel_mrays_per_s": r["kernel_mrays_per_s"]}))'
done; done
