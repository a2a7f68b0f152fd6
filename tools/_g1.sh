set -x
export SSEGPU_ORDERED_SINGLE=1
timeout 900 python -m pytest tests/test_gpu_frame.py -m gpu -x -q -k "ordered or op or full_size or bit_exact or boundaries or sparse" > gpurun_out/r02_gputest_o4.log 2>&1; tail -5 gpurun_out/r02_gputest_o4.log
for w in ordered; do timeout 200 python tools/prof_target.py $w > gpurun_out/r02_o4_$w.log 2>&1; tail -2 gpurun_out/r02_o4_$w.log; done
timeout 200 python tools/time_kernel.py --op 8 --noncanon 0.1 2>&1 | tail -1
