timeout 1200 python -m pytest tests/test_gpu_frame.py tests/test_gpu_inspect.py tests/test_gpu_ext_modes.py tests/test_gpu_interest.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
for w in plain plain32 inspect_ab match bridge ordered stream; do timeout 200 python tools/prof_target.py $w 2>&1 | tail -1; done
