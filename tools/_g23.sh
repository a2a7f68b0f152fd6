set -x
for w in inspect_ab match; do timeout 120 python tools/prof_target.py $w > gpurun_out/r02_t11_$w.log 2>&1; tail -1 gpurun_out/r02_t11_$w.log; done
timeout 600 python -m pytest tests/test_gpu_inspect.py tests/test_gpu_ext_modes.py -m gpu -x -q 2>&1 | tail -3
