timeout 1200 python tests/soak_fuzz.py --seeds 24 --records 200000 --first-seed 200 > gpurun_out/r02_soak3.log 2>&1; tail -2 gpurun_out/r02_soak3.log
