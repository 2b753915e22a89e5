for G in 6500 5500 4500 3500; do
  echo -n "hyq LANES=2 GCAP=$G " >> gpurun_out/hyq.log
  TOWR_B200_LANES=2 TOWR_B200_GCAP=$G TOWR_B200_DEBUG=1 timeout 300 python bench.py --workload hyq_gallop_durations --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 1 2>&1 | python -c "
import sys, json
ls = sys.stdin.read().strip().split('\n')
info = [l for l in ls if l.startswith('batch:')]
try:
    j = json.loads(ls[-1]); print(j['ms_per_step'], round(j['roofline']['frac'], 4), ' | '.join(info[-2:]))
except Exception as e:
    print('FAIL', ls[-1][-200:])
" >> gpurun_out/hyq.log
done
