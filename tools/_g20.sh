set -x
timeout 1500 python tests/soak_fuzz.py --seeds 6 --records 100000 > gpurun_out/r02_soak.log 2>&1; tail -12 gpurun_out/r02_soak.log
