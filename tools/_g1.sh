timeout 900 python -m pytest tests/test_gpu_interest.py tests/test_gpu_churn.py tests/test_gpu_ext_modes.py -m gpu -x -q 2>&1 | tail -2
for w in match bridge; do timeout 200 python tools/prof_target.py $w 2>&1 | tail -1; done
