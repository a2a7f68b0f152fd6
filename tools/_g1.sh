timeout 1500 python -m pytest tests/test_gpu_frame.py tests/test_gpu_ext_modes.py tests/test_zz_no_alloc.py tests/test_gpu_adapter_host.py -m gpu -x -q 2>&1 | tail -2
for w in plain ordered; do timeout 200 python tools/prof_target.py $w 2>&1 | tail -1; done
for p in 0.1 1; do timeout 200 python tools/time_kernel.py --op 8 --noncanon $p 2>&1 | tail -1; done
timeout 200 python tools/time_kernel.py --op 9 2>&1 | tail -1
timeout 200 python tools/time_kernel.py --op 137 2>&1 | tail -1
python tools/prof_target.py ordered > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 16 -c 8 --csv --log-file gpurun_out/r02_launches_ordered.csv python tools/prof_target.py ordered > gpurun_out/ncu1.log 2>&1
