set -x
mkdir -p gpurun_out
B="python bench.py --steps 3 --warmup 3 --e2e-steps 1 --no-cpu-baseline --no-extra"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r30_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r30_tests.log
timeout 400 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r30_bench_ref.json 2> gpurun_out/r30_bench_ref.err
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r30_bench.json 2> gpurun_out/r30_bench.err
timeout 300 python tools/coalescer_bench.py gpurun_out/r30_coalescer.json > gpurun_out/r30_coalescer.log 2>&1
for w in anymal_trot_durations biped_walk_durations hyq_gallop_durations monoped_hop; do timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extra --workload $w >> gpurun_out/r30_workloads.jsonl 2>> gpurun_out/r30_workloads.err; done
timeout 300 python tools/trajectory_bench.py > gpurun_out/r30_trajectory.log 2>&1
timeout 600 python tools/parity_survey.py gpu > gpurun_out/r30_parity_survey_gpu.txt 2>&1
timeout 900 python tools/mixed_sweep.py gpurun_out/r30_mixed.json > gpurun_out/r30_mixed.log 2>&1
timeout 300 $B > gpurun_out/r30_plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r30_launches.csv $B > gpurun_out/r30_ncu1.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:EvalKernel -s 3 -c 1 -f -o gpurun_out/r30_prof $B > gpurun_out/r30_ncu2.log 2>&1
tail -n 3 gpurun_out/r30_tests.log; cat gpurun_out/r30_bench.json
