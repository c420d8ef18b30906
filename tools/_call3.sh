mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -2 | tee gpurun_out/r02_gpu_tests.log
timeout 1200 python bench.py > gpurun_out/r02_bench_default.json 2> gpurun_out/r02_bench_default.err; tail -c 300 gpurun_out/r02_bench_default.err
python -c "
import json
d=json.load(open('gpurun_out/r02_bench_default.json'))
print('value',d['value'],'frac',d['roofline']['frac'],'e2e',d['e2e']['value'])
for o in d['roofline']['others']: print(o['workload'],o['frac'])
"
