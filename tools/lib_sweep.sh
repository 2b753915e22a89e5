#!/bin/bash
# bench (device-resident value only) with alternative builds of the library over workloads:
#   tools/lib_sweep.sh "<lib suffix list, base = default>" "<workload list>"
mkdir -p gpurun_out
for W in $2; do for S in $1; do
  [ "$S" = base ] && S=""
  echo -n "$W lib=$S " >> gpurun_out/libsweep.log
  TOWR_B200_LIB=$PWD/towr2025_b200/libtowr_b200$S.so timeout 300 python bench.py --workload $W --steps 30 --warmup 5 --no-cpu-baseline --e2e-steps 1 2>&1 | tail -1 | python -c "
import sys, json
l = sys.stdin.read().strip()
try:
    j = json.loads(l); print(j['ms_per_step'], round(j['roofline']['frac'], 4))
except Exception as e:
    print('FAIL', l[-200:])
" >> gpurun_out/libsweep.log
done; done
cat gpurun_out/libsweep.log
