mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/quick_tests.log 2>&1; tail -n 3 gpurun_out/quick_tests.log
for i in 1 2; do timeout 120 python bench.py --steps 20 --warmup 5 --e2e-steps 3 --no-cpu-baseline --no-extra 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('bench', round(d['ms_per_step'], 4), round(d['roofline']['frac'], 4), round(d['e2e']['ms_per_step'], 2))
"; done
