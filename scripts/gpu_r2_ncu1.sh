mkdir -p gpurun_out
B="./bin/gol 6 16384 128 256 0 --k 4 --mode persistent"
GOL_FLOW_FLAGS=0 $B > gpurun_out/plain_a.log 2>&1 && \
GOL_FLOW_FLAGS=0 ncu --set full --clock-control none --import-source on -k regex:life_flow_kernel -c 1 -o gpurun_out/prof_flow_owner $B > gpurun_out/ncu_a.log 2>&1
tail -2 gpurun_out/ncu_a.log
GOL_FLOW_FLAGS=4 $B --strip 31 > gpurun_out/plain_b.log 2>&1 && \
GOL_FLOW_FLAGS=4 ncu --set full --clock-control none --import-source on -k regex:life_flow_kernel -c 1 -o gpurun_out/prof_flow_ticket31 $B --strip 31 > gpurun_out/ncu_b.log 2>&1
tail -2 gpurun_out/ncu_b.log
GOL_FLOW_FLAGS=5 $B --strip 31 > gpurun_out/plain_c.log 2>&1 && \
GOL_FLOW_FLAGS=5 ncu --set full --clock-control none --import-source on -k regex:life_flow_kernel -c 1 -o gpurun_out/prof_flow_ticket31_nodeps $B --strip 31 > gpurun_out/ncu_c.log 2>&1
tail -2 gpurun_out/ncu_c.log
ls -la gpurun_out/*.ncu-rep
