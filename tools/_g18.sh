set -x
timeout 600 python -m pytest tests/test_gpu_chunk.py tests/test_gpu_wire.py tests/test_gpu_encode.py -m gpu -x -q > gpurun_out/r02_gputest11.log 2>&1; tail -15 gpurun_out/r02_gputest11.log
