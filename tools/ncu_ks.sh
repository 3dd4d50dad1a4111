python tools/probe_ks.py 16384 0 > gpurun_out/plain_ks.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_keyswitch_step -s 2 -c 1 -f -o gpurun_out/r2_prof_ks python tools/probe_ks.py 16384 0 > gpurun_out/ncu_ks.log 2>&1
echo rc=$?
