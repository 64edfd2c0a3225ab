mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
./bin/gol 6 16384 64 256 0 --k 4 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:life_fused_kernel -s 6 -c 1 -o gpurun_out/prof_fused_k4_final ./bin/gol 6 16384 64 256 0 --k 4 > gpurun_out/ncu1.log 2>&1
tail -1 gpurun_out/ncu1.log
./bin/gol 6 16384 64 256 0 --k 4 --mode persistent > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:life_persistent -c 1 -o gpurun_out/prof_persistent_k4 ./bin/gol 6 16384 64 256 0 --k 4 --mode persistent > gpurun_out/ncu2.log 2>&1
tail -1 gpurun_out/ncu2.log
./bin/gol 6 16384 8 256 0 --mode generic > gpurun_out/plain3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv ./bin/gol 6 16384 16 256 0 --k 4 > gpurun_out/ncu3.log 2>&1
tail -3 gpurun_out/launches.csv
