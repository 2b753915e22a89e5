#!/bin/bash
# sweep of the kernel's build-time / run-time knobs on a B200 (device-resident value only)
# usage: tools/sweep.sh "<lanes list>" "<gcap list>"
mkdir -p gpurun_out
for L in $1; do for G in $2; do
  echo -n "LANES=$L GCAP=$G " >> gpurun_out/sweep.log
  TOWR_B200_LANES=$L TOWR_B200_GCAP=$G timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 1 2>&1 | tail -1 | python -c "
import sys, json
l = sys.stdin.read().strip()
try:
    j = json.loads(l); print(j['ms_per_step'], j['roofline']['frac'])
except Exception as e:
    print('FAIL', l[-200:])
" >> gpurun_out/sweep.log
done; done
cat gpurun_out/sweep.log
