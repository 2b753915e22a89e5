# Smoke run of every kernel family (default, extended instantiations, multi-part classes, dynamic-only transfer, samplers).
# compute-sanitizer is closed on this pool ("runs under it have left GPUs needing a reset"), so this is the plain run; on
# a machine where it is open:  compute-sanitizer --tool memcheck python tools/sanitize_smoke.py
mkdir -p gpurun_out
timeout 300 python tools/sanitize_smoke.py > gpurun_out/sanitize_plain.log 2>&1; echo "rc=$?"; tail -n 8 gpurun_out/sanitize_plain.log | cut -c1-160
