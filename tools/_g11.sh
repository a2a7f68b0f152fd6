set -x
timeout 900 python -m pytest tests/test_gpu_inspect.py tests/test_gpu_ext_modes.py -m gpu -x -q > gpurun_out/r02_gputest6.log 2>&1; tail -5 gpurun_out/r02_gputest6.log
for w in stream; do timeout 120 python tools/prof_target.py $w > gpurun_out/r02_t7_$w.log 2>&1; tail -1 gpurun_out/r02_t7_$w.log; done
python tools/prof_target.py stream > gpurun_out/plain1.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 14 --csv --log-file gpurun_out/r02_launches_stream_v5.csv python tools/prof_target.py stream > gpurun_out/ncu1.log 2>&1
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_v1.json 2> gpurun_out/r02_bench_v1.err; tail -5 gpurun_out/r02_bench_v1.err; cat gpurun_out/r02_bench_v1.json | cut -c1-3000
timeout 600 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/r02_bench_v1_ref.json 2>> gpurun_out/r02_bench_v1.err; cat gpurun_out/r02_bench_v1_ref.json | cut -c1-600
