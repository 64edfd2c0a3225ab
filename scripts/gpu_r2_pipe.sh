mkdir -p gpurun_out; : > gpurun_out/pipe2.log
timeout 120 python scripts/pipe_exp2.py auto >> gpurun_out/pipe2.log 2>&1
./bin/gol 6 16384 1024 256 0 --json --k 4 | tail -1 | cut -c1-100 >> gpurun_out/pipe2.log
./bin/gol 6 65536 64 256 0 --json --k 4 | tail -1 | cut -c1-100 >> gpurun_out/pipe2.log
cat gpurun_out/pipe2.log
