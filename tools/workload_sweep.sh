for W in anymal_trot_durations hyq_gallop_durations biped_walk_durations monoped_hop; do for K in lockstep bulk; do
  echo -n "$W $K " >> gpurun_out/rt.log
  TOWR_B200_KERNEL=$K TOWR_B200_DEBUG=1 timeout 300 python bench.py --workload $W --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 1 2>&1 | python -c "
import sys, json
ls = sys.stdin.read().strip().split('\n')
info = [l for l in ls if l.startswith('batch:')]
try:
    j = json.loads(ls[-1]); print(j['ms_per_step'], round(j['roofline']['frac'], 4), j['value'], ' | '.join(info[-2:]))
except Exception as e:
    print('FAIL', ls[-1][-200:])
" >> gpurun_out/rt.log
done; done
