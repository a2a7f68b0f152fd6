set -x
timeout 600 python -m pytest tests/test_gpu_wire.py tests/test_gpu_frame.py -m gpu -x -q > gpurun_out/r02_gputest10.log 2>&1; tail -8 gpurun_out/r02_gputest10.log
