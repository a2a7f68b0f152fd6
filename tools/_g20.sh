set -x
timeout 1500 python tests/soak_fuzz.py --seeds 24 --records 200000 > gpurun_out/r02_soak.log 2>&1; tail -6 gpurun_out/r02_soak.log
