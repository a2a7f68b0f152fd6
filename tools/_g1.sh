timeout 1500 python -m pytest tests/test_gpu_frame.py tests/test_gpu_inspect.py tests/test_gpu_ext_modes.py tests/test_gpu_interest.py tests/test_zz_no_alloc.py tests/test_gpu_adapter_host.py -m gpu -x -q 2>&1 | tail -2
for w in plain plain32 inspect_ab match bridge ordered stream; do timeout 200 python tools/prof_target.py $w 2>&1 | tail -1; done
for p in 0.1 1; do timeout 200 python tools/time_kernel.py --op 0 --noncanon $p 2>&1 | tail -1; timeout 200 python tools/time_kernel.py --op 8 --noncanon $p 2>&1 | tail -1; done
