set -x
timeout 900 python -m pytest tests/test_gpu_inspect.py tests/test_gpu_ext_modes.py -m gpu -x -q > gpurun_out/r02_gputest_i3.log 2>&1; tail -3 gpurun_out/r02_gputest_i3.log
for w in inspect inspect_ab; do timeout 200 python tools/prof_target.py $w > gpurun_out/r02_i3_$w.log 2>&1; tail -1 gpurun_out/r02_i3_$w.log; done
