set -x
export SSEGPU_LIB=$PWD/tools/exp/libssegpu_B.so
for p in 1 10 100; do timeout 200 python tools/time_kernel.py --op 0 --noncanon $p 2>&1 | tail -1; done
