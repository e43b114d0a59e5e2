# round-1 closing evidence on one GPU (final code): default bench, reference arm, smoke, GPU tests
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py > gpurun_out/r01l_bench_full.json 2> gpurun_out/r38_bench.err; echo "bench rc=$?"
python bench.py --impl reference > gpurun_out/r01l_bench_reference.json 2> gpurun_out/r38_ref.err; echo "ref rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r01l_bench_full.json')); r=d['roofline']
print(round(d['value'],1), round(d['ms_per_step'],1), round(d['mrays_per_s'],1), 'e2e', round(d['e2e']['value'],1), 'frac', round(r['frac'],3), r['l2'])
d=json.load(open('gpurun_out/r01l_bench_reference.json')); print('ref', d['value'], d['cpu_baseline']['cores'])"
