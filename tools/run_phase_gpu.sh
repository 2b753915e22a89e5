mkdir -p gpurun_out
TIMING_CODE=2110 TOWR_B200_DEBUG=1 timeout 300 python tools/phase_timing.py 4096 durations > gpurun_out/phase_2110.log 2>&1; tail -n 25 gpurun_out/phase_2110.log | cut -c1-220
