set -x
timeout 900 python -m pytest tests/test_gpu_churn.py tests/test_gpu_ext_modes.py tests/test_gpu_interest.py -m gpu -x -q -s > gpurun_out/r02_gputest7.log 2>&1; tail -15 gpurun_out/r02_gputest7.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputest8.log 2>&1; tail -5 gpurun_out/r02_gputest8.log
for w in stream match; do timeout 120 python tools/prof_target.py $w > gpurun_out/r02_t8_$w.log 2>&1; tail -1 gpurun_out/r02_t8_$w.log; done
