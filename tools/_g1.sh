set -x
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputest_full2.log 2>&1; tail -4 gpurun_out/r02_gputest_full2.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r02_bench_v2.json 2> gpurun_out/r02_bench_v2.err; tail -c 600 gpurun_out/r02_bench_v2.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r02_bench_v2_reference_arm.json 2>/dev/null; tail -c 400 gpurun_out/r02_bench_v2_reference_arm.json
