set -x
python tools/prof_target.py stream > gpurun_out/plain1.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_stream_walk -s 3 -c 1 -o gpurun_out/r02_prof_walk_v2 python tools/prof_target.py stream > gpurun_out/ncu1.log 2>&1
python tools/prof_target.py inspect_ab > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 4 -c 1 -o gpurun_out/r02_prof_inspect_v2 python tools/prof_target.py inspect_ab > gpurun_out/ncu2.log 2>&1
