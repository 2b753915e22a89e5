mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/rt_tests.log 2>&1; tail -n 3 gpurun_out/rt_tests.log
for w in anymal_trot anymal_trot_durations biped_walk_durations hyq_gallop_durations; do timeout 200 python bench.py --steps 10 --warmup 3 --e2e-steps 1 --no-cpu-baseline --no-extra --workload $w 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$w', round(d['ms_per_step'], 4), round(d['roofline']['frac'], 4))
"; done
