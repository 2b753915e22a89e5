#!/bin/bash
# bench (device-resident value only) over the kernel variants and stream skews on a B200
# usage: tools/variant_sweep.sh "<workload list>" "<variant[:skew] list>"       e.g. "anymal_trot" "lockstep ring streams:12000"
mkdir -p gpurun_out
for W in $1; do for V in $2; do
  K=${V%%:*}; SK=0; [ "$K" != "$V" ] && SK=${V##*:}
  echo -n "$W KERNEL=$K SKEW=$SK " >> gpurun_out/variants.log
  TOWR_B200_KERNEL=$K TOWR_B200_SKEW=$SK TOWR_B200_DEBUG=1 timeout 300 python bench.py --workload $W --steps 50 --warmup 5 --no-cpu-baseline --e2e-steps 1 2>&1 | python -c "
import sys, json
ls = sys.stdin.read().strip().split('\n')
info = [l for l in ls if l.startswith('batch:')]
try:
    j = json.loads(ls[-1]); print(j['ms_per_step'], round(j['roofline']['frac'], 4), info[-1] if info else '')
except Exception as e:
    print('FAIL', ls[-1][-300:])
" >> gpurun_out/variants.log
done; done
cat gpurun_out/variants.log
