# TOWR_B200_KERNEL=split against the default kernel: parity tests, then bench.py per workload.  Short timeouts: a
# producer / consumer hand-off that deadlocks must not eat the GPU budget.
mkdir -p gpurun_out
B="python bench.py --steps 20 --warmup 5 --e2e-steps 1 --no-cpu-baseline --no-extra"
show() { python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$1', round(d['ms_per_step'], 4), round(d['roofline']['frac'], 4), d['roofline']['kernel'])
    elif 'rror' in l: print('$1', l.strip()[:300])
"; }
TOWR_B200_KERNEL=split timeout 90 $B 2>&1 | show "anymal_trot split"
if [ ${PIPESTATUS[0]} -ne 0 ]; then echo "split bench failed or hung: stopping"; exit 1; fi
TOWR_B200_KERNEL=lockstep timeout 90 $B 2>&1 | show "anymal_trot lockstep"
timeout 300 python -m pytest tests/test_kernel_variants.py -x -q -k split > gpurun_out/split_tests.log 2>&1; tail -n 5 gpurun_out/split_tests.log
for w in anymal_trot_durations biped_walk_durations hyq_gallop_durations monoped_hop; do
  for k in lockstep split; do TOWR_B200_KERNEL=$k timeout 90 $B --workload $w 2>&1 | show "$w $k"; done
done
for cap in 3000 3400 3800; do TOWR_B200_GCAP=$cap TOWR_B200_KERNEL=split timeout 90 $B 2>&1 | show "anymal_trot split gcap $cap"; done
