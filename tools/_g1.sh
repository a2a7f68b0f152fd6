timeout 1200 python -m pytest tests/test_gpu_frame.py tests/test_gpu_encode.py -m gpu -x -q 2>&1 | tail -2
for w in plain plain32 inspect_ab match bridge ordered; do timeout 200 python tools/prof_target.py $w 2>&1 | tail -1; done
timeout 300 python tools/bench_encode.py 2>&1 | tail -2
