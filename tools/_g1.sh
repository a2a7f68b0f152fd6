set -x
timeout 900 python -m pytest tests/test_gpu_frame.py tests/test_gpu_ext_modes.py -m gpu -x -q > gpurun_out/r02_gputest_o5.log 2>&1; tail -5 gpurun_out/r02_gputest_o5.log
for w in ordered; do timeout 200 python tools/prof_target.py $w > gpurun_out/r02_o5_$w.log 2>&1; tail -1 gpurun_out/r02_o5_$w.log; done
for p in 0.1 1 10 100; do timeout 200 python tools/time_kernel.py --op 8 --noncanon $p 2>&1 | tail -1; done
for p in 0.1 1 10 100; do timeout 200 python tools/time_kernel.py --op 0 --noncanon $p 2>&1 | tail -1; done
