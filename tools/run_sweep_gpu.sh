mkdir -p gpurun_out
timeout 1500 python tools/mixed_sweep.py gpurun_out/r18_mixed.json > gpurun_out/r18_mixed.log 2>&1; tail -n 3 gpurun_out/r18_mixed.log | cut -c1-250
