set -x
go version > gpurun_out/go_probe.txt 2>&1; nproc >> gpurun_out/go_probe.txt; nvidia-smi topo -m >> gpurun_out/go_probe.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputest0.log 2>&1; tail -3 gpurun_out/r02_gputest0.log
for w in inspect_ab match bridge ordered; do python tools/prof_target.py $w > gpurun_out/r02_t0_$w.log 2>&1; cat gpurun_out/r02_t0_$w.log | tail -1; done
python tools/prof_target.py inspect_ab > gpurun_out/plain1.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 4 -c 1 -o gpurun_out/r02_prof_inspect_v0 python tools/prof_target.py inspect_ab > gpurun_out/ncu1.log 2>&1
python tools/prof_target.py match > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 4 -c 1 -o gpurun_out/r02_prof_match_v0 python tools/prof_target.py match > gpurun_out/ncu2.log 2>&1
python tools/prof_target.py ordered > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 8 -c 2 -o gpurun_out/r02_prof_ordered_v0 python tools/prof_target.py ordered > gpurun_out/ncu3.log 2>&1
ls -la gpurun_out
