mkdir -p gpurun_out
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_ours.log 2>&1; echo "bench exit $?"
tail -1 gpurun_out/bench_ours.log | cut -c1-200
GOL_REF_BUDGET_S=60 timeout 600 python bench.py --impl reference --steps 6 --warmup 3 > gpurun_out/bench_ref.log 2>&1; echo "ref exit $?"
tail -1 gpurun_out/bench_ref.log | cut -c1-120
./bin/gol 6 16384 64 256 0 --k 4 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:life_fused_kernel -s 6 -c 1 -o gpurun_out/prof_fused_k4_final ./bin/gol 6 16384 64 256 0 --k 4 > gpurun_out/ncu1.log 2>&1
tail -1 gpurun_out/ncu1.log
./bin/gol 6 65536 16 256 0 --k 4 > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none -k regex:life_fused_kernel -s 2 -c 1 -o gpurun_out/prof_fused_k4_65536 ./bin/gol 6 65536 16 256 0 --k 4 > gpurun_out/ncu2.log 2>&1
tail -1 gpurun_out/ncu2.log
python __graft_entry__.py smoke 2>&1 | tail -1
