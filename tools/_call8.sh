mkdir -p gpurun_out
( nvidia-smi topo -m; lscpu | grep -E "^CPU\(s\)|NUMA|Socket|Model name"; free -g | head -2; for d in /sys/bus/pci/devices/*; do if [ -f $d/class ] && grep -q 0x0302 $d/class; then echo $d $(cat $d/numa_node) $(cat $d/local_cpulist); fi; done ) > gpurun_out/r02_box_topology_8gpu.txt 2>&1
python tools/pcie_ceiling.py --gpus 8 --iters 10 > gpurun_out/r02_pcie_8.json 2> gpurun_out/r02_pcie_8.err
python tools/pcie_ceiling.py --gpus 4 --iters 10 > gpurun_out/r02_pcie_4.json 2> gpurun_out/r02_pcie_4.err
cat gpurun_out/r02_pcie_8.json gpurun_out/r02_pcie_4.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 50 --warmup 3 --no-sub > gpurun_out/r02_bench_8gpu_nosub.json 2> gpurun_out/r02_bench_8.err
python -c "
import json
d=json.load(open('gpurun_out/r02_bench_8gpu_nosub.json'))
print('value',d['value'],'frac',d['roofline']['frac'],'e2e',d['e2e'])
"
head -c 1500 gpurun_out/r02_box_topology_8gpu.txt
