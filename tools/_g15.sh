set -x
python tools/pcie_probe.py --gpus 8 > gpurun_out/r02_pcie_probe_8gpu.txt 2>&1; head -1 gpurun_out/r02_pcie_probe_8gpu.txt | cut -c1-1800
python tools/pcie_probe.py --gpus 4 > gpurun_out/r02_pcie_probe_4gpu.txt 2>&1; head -1 gpurun_out/r02_pcie_probe_4gpu.txt | cut -c1-200
