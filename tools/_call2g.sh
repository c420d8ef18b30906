mkdir -p gpurun_out
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02_bench_2gpu.json 2> gpurun_out/r02_bench_2gpu.err ) 2>&1 | tail -3
tail -c 600 gpurun_out/r02_bench_2gpu.err
python -c "
import json
d=json.load(open('gpurun_out/r02_bench_2gpu.json'))
print('value',d['value'],'frac',d['roofline']['frac'],'e2e',d['e2e']['value'],d['e2e']['frac_of_bidirectional_ceiling'])
for o in d['roofline']['others']: print(o['workload'],o['kernel'],o['achieved'],o['frac'])
print(d['sub']['e2e_32_streams_256MiB_blocks'])
"
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --impl reference --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02_bench_2gpu_ref.json ) 2>&1 | tail -3
python -c "
import json
d=json.load(open('gpurun_out/r02_bench_2gpu_ref.json')); print(d['value'], d['steps'], d['warmup'], d['cpu_baseline']['cores'])"
