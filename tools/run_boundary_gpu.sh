mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r16_tests.log 2>&1; tail -n 6 gpurun_out/r16_tests.log
timeout 120 python bench.py --steps 20 --warmup 5 --e2e-steps 2 --no-cpu-baseline --no-extra 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('bench', round(d['ms_per_step'], 4), round(d['roofline']['frac'], 4), round(d['e2e']['ms_per_step'], 2))
"
timeout 900 python tools/mixed_sweep.py gpurun_out/r16_mixed.json > gpurun_out/r16_mixed.log 2>&1; tail -n 2 gpurun_out/r16_mixed.log | cut -c1-200
