set -x
timeout 600 python -m pytest tests/test_gpu_ext_modes.py tests/test_gpu_churn.py -m gpu -x -q > gpurun_out/r02_gputest12.log 2>&1; tail -15 gpurun_out/r02_gputest12.log
