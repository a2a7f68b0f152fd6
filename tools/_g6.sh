set -x
timeout 900 python -m pytest tests/test_gpu_inspect.py tests/test_gpu_ext_modes.py tests/test_gpu_interest.py -m gpu -x -q > gpurun_out/r02_gputest3.log 2>&1; tail -5 gpurun_out/r02_gputest3.log
for w in match stream inspect_ab; do timeout 120 python tools/prof_target.py $w > gpurun_out/r02_t4_$w.log 2>&1; tail -1 gpurun_out/r02_t4_$w.log; done
python tools/prof_target.py stream > gpurun_out/plain1.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 14 --csv --log-file gpurun_out/r02_launches_stream_v2.csv python tools/prof_target.py stream > gpurun_out/ncu1.log 2>&1
