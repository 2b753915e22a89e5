mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --e2e-steps 1 --no-cpu-baseline --no-extra --workload anymal_trot_durations"
timeout 300 $B > gpurun_out/rt_plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:EvalKernel -s 3 -c 1 -f -o gpurun_out/rt_prof $B > gpurun_out/rt_ncu.log 2>&1
tail -n 2 gpurun_out/rt_ncu.log | cut -c1-200
