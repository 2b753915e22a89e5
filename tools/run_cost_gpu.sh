# GPU check of the cost entry points (NodeCost, EnergyCost, AngularMomentumCost; batch, device and static variants)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cost.py tests/test_zz_no_alloc.py tests/test_hopper_example.py tests/test_biped_example.py -m gpu -x -q > gpurun_out/cost_tests.log 2>&1; tail -n 15 gpurun_out/cost_tests.log
timeout 120 python tools/cost_timing.py > gpurun_out/cost_timing.json 2> gpurun_out/cost_timing.err; cat gpurun_out/cost_timing.json; tail -n 3 gpurun_out/cost_timing.err
