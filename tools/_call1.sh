set -x
mkdir -p gpurun_out
( which cargo rustc rustup; ls -d ~/.cargo ~/.rustup /usr/local/cargo /opt/rust* 2>&1; echo "cargo probe rc=$?" ) > gpurun_out/r02_toolchain_probe.txt 2>&1
( nvidia-smi topo -m; lscpu | head -30; ls /sys/devices/system/node/; for d in /sys/bus/pci/devices/*; do if [ -f $d/class ] && grep -q 0x0302 $d/class; then echo $d $(cat $d/numa_node) $(cat $d/local_cpulist); fi; done; nproc; free -g ) > gpurun_out/r02_box_topology.txt 2>&1
python -m pytest tests/test_gpu_parity.py -x -q -k "two_devices" > gpurun_out/r02_two_devices.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_two_devices.log
python tools/pcie_ceiling.py --gpus 1 > gpurun_out/r02_pcie_1.json 2> gpurun_out/r02_pcie_1.err
python tools/pcie_ceiling.py --gpus 2 > gpurun_out/r02_pcie_2.json 2> gpurun_out/r02_pcie_2.err
tail -3 gpurun_out/r02_two_devices.log; cat gpurun_out/r02_pcie_1.json gpurun_out/r02_pcie_2.json; cat gpurun_out/r02_toolchain_probe.txt
