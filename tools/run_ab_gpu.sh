mkdir -p gpurun_out; rm -f gpurun_out/libsweep.log
bash tools/lib_sweep.sh "base _r12 _nohard base _r12 _nohard" "anymal_trot" > /dev/null 2>&1
cat gpurun_out/libsweep.log
