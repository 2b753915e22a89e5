#!/bin/bash
# sweep of the kernel's run-time knobs on a B200 (device-resident value only)
# usage: tools/sweep.sh "<lanes list>" "<gcap list>" ["<kernel list>"]     (gcap "-" = the builder's default)
mkdir -p gpurun_out
for K in ${3:-default}; do for L in $1; do for G in $2; do
  echo -n "KERNEL=$K LANES=$L GCAP=$G " >> gpurun_out/sweep.log
  ( [ "$G" != "-" ] && export TOWR_B200_GCAP=$G; [ "$K" != default ] && export TOWR_B200_KERNEL=$K
  TOWR_B200_DEBUG=1 TOWR_B200_LANES=$L timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 1 2>&1 | python -c "
import sys, json
ls = sys.stdin.read().strip().split('\n')
info = [l for l in ls if l.startswith('batch:')]
try:
    j = json.loads(ls[-1]); print(j['ms_per_step'], round(j['roofline']['frac'], 4), info[-1] if info else '')
except Exception as e:
    print('FAIL', ls[-1][-200:])
" ) >> gpurun_out/sweep.log
done; done; done
cat gpurun_out/sweep.log
