timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputest_full3.log 2>&1; tail -3 gpurun_out/r02_gputest_full3.log
timeout 900 python tests/soak_fuzz.py --seeds 24 --records 200000 --first-seed 100 > gpurun_out/r02_soak2.log 2>&1; tail -2 gpurun_out/r02_soak2.log
timeout 900 python bench.py > gpurun_out/r02_bench_v3.json 2> gpurun_out/r02_bench_v3.err; tail -c 300 gpurun_out/r02_bench_v3.json
