#!/bin/bash
# bench of library variants of the warp-specialised kernels: tools/ws_split_sweep.sh "<lib suffixes>" <kernel>
for S in $1; do [ "$S" = base ] && S=""
  echo -n "$2 lib=$S " >> gpurun_out/ws.log
  TOWR_B200_LIB=$PWD/towr2025_b200/libtowr_b200$S.so TOWR_B200_KERNEL=$2 timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 1 2>&1 | tail -1 | python -c "
import sys, json
l = sys.stdin.read().strip()
try:
    j = json.loads(l); print(j['ms_per_step'], round(j['roofline']['frac'], 4))
except Exception as e:
    print('FAIL', l[-200:])
" >> gpurun_out/ws.log
done
cat gpurun_out/ws.log
