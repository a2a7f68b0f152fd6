timeout 900 python -m pytest tests/test_gpu_frame.py -m gpu -x -q -k "alignment or boundaries" 2>&1 | tail -2
timeout 900 python -m pytest tests/test_gpu_interest.py tests/test_gpu_churn.py -m gpu -x -q 2>&1 | tail -2
for w in match inspect_ab; do timeout 200 python tools/prof_target.py $w 2>&1 | tail -1; done
