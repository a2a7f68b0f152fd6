set -x
python tools/prof_target.py match > gpurun_out/plain1.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 4 -c 1 -o gpurun_out/r02_prof_match_v3 python tools/prof_target.py match > gpurun_out/ncu1.log 2>&1
python tools/prof_target.py bridge > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 4 -c 1 -o gpurun_out/r02_prof_bridge_v3 python tools/prof_target.py bridge > gpurun_out/ncu2.log 2>&1
