mkdir -p gpurun_out
( time timeout 1200 python bench.py > gpurun_out/r02_bench_default.json 2> gpurun_out/r02_bench_default.err ) 2>&1 | tail -3
tail -c 600 gpurun_out/r02_bench_default.err
python -c "
import json
d=json.load(open('gpurun_out/r02_bench_default.json'))
print('value',d['value'],'frac',d['roofline']['frac'],'traffic',d['roofline']['traffic'],'e2e',d['e2e'])
for o in d['roofline']['others']: print(o)
print(d['clocks']); print(d['cpu_baseline']['value'], d['cpu_baseline']['single_thread_value'])
print(d['sub']['e2e_32_streams_256MiB_blocks']); print(d['sub']['small_call_latency_us'])
"
timeout 300 python tools/exp_kernels.py fused 128 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:rs_x2_crc_fast -s 2 -c 1 -o gpurun_out/r02_fused_final_prof python tools/exp_kernels.py fused 128 > gpurun_out/ncu2.log 2>&1
cat gpurun_out/plain.log
