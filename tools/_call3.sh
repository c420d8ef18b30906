mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r02_gpu_tests.log
( time timeout 1200 python bench.py > gpurun_out/r02_bench_default.json 2> gpurun_out/r02_bench_default.err ) 2>&1 | tail -3
tail -c 400 gpurun_out/r02_bench_default.err
python -c "
import json
d=json.load(open('gpurun_out/r02_bench_default.json'))
print('value',d['value'],'frac',d['roofline']['frac'],'e2e',d['e2e']['value'],d['e2e']['frac_of_h2d_ceiling'])
for o in d['roofline']['others']: print(o['workload'],o['kernel'],o['frac'],o.get('cpu_port_GBps'))
print(d['sub']['block_checksums_batch'])
"
timeout 300 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"rs_lut8_kernel|crc_segments|rs_x2_crc_fast|crc_sfold_finish" -c 12 -o gpurun_out/r02_bench_prof python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu2.log 2>&1
tail -2 gpurun_out/ncu2.log
