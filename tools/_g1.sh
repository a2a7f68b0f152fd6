set -x
timeout 1500 python -m pytest tests/test_gpu_frame.py tests/test_gpu_ext_modes.py -m gpu -x -q > gpurun_out/r02_gputest_o2.log 2>&1; tail -5 gpurun_out/r02_gputest_o2.log
for w in ordered inspect_ab; do timeout 200 python tools/prof_target.py $w > gpurun_out/r02_o2_$w.log 2>&1; tail -2 gpurun_out/r02_o2_$w.log; done
python tools/prof_target.py ordered > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 16 -c 8 --csv --log-file gpurun_out/r02_launches_ordered_o2.csv python tools/prof_target.py ordered > gpurun_out/ncu1.log 2>&1
