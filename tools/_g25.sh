set -x
P=gpurun_out
python tools/prof_target.py plain > $P/p.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 4 -c 1 -o $P/r02_ncu_plain python tools/prof_target.py plain > $P/n.log 2>&1
python tools/prof_target.py ordered > $P/p.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 8 -c 2 -o $P/r02_ncu_ordered python tools/prof_target.py ordered > $P/n.log 2>&1
python tools/prof_target.py inspect_ab > $P/p.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_frame_tiles -s 4 -c 1 -o $P/r02_ncu_inspect python tools/prof_target.py inspect_ab > $P/n.log 2>&1
ls -la $P/*.ncu-rep
