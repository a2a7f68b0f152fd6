set -x
python tools/pcie_probe.py --gpus 2 > gpurun_out/r02_pcie_probe_2gpu.txt 2>&1; head -3 gpurun_out/r02_pcie_probe_2gpu.txt | cut -c1-600
