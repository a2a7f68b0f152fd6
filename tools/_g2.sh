set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputest1.log 2>&1; tail -30 gpurun_out/r02_gputest1.log
for w in inspect_ab match bridge plain; do timeout 120 python tools/prof_target.py $w > gpurun_out/r02_t1_$w.log 2>&1; tail -1 gpurun_out/r02_t1_$w.log; done
