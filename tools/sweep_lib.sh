#!/bin/bash
# bench (device-resident value only) with alternative builds of the library: tools/sweep_lib.sh "<suffix list>" "<lanes>" "<gcap>"
mkdir -p gpurun_out
for S in $1; do [ "$S" = base ] && S=""; for L in $2; do for G in $3; do
  echo -n "LIB=$S LANES=$L GCAP=$G " >> gpurun_out/sweep.log
  TOWR_B200_LIB=$PWD/towr2025_b200/libtowr_b200$S.so TOWR_B200_LANES=$L TOWR_B200_GCAP=$G timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 1 2>&1 | tail -1 | python -c "
import sys, json
l = sys.stdin.read().strip()
try:
    j = json.loads(l); print(j['ms_per_step'], j['roofline']['frac'])
except Exception as e:
    print('FAIL', l[-200:])
" >> gpurun_out/sweep.log
done; done; done
tail -n 12 gpurun_out/sweep.log
