mkdir -p gpurun_out; : > gpurun_out/strip.log
for st in 52 58 63 66 72 84 126; do timeout 60 ./bin/gol 6 16384 1024 256 0 --json --k 4 --strip $st 2>&1 | tail -1 >> gpurun_out/strip.log; done
for st in 63; do GOL_PDL=1 timeout 60 ./bin/gol 6 16384 1024 256 0 --json --k 4 --strip $st 2>&1 | tail -1 >> gpurun_out/strip.log; done
timeout 60 ./bin/gol 6 16384 1024 256 0 --json --k 3 2>&1 | tail -1 >> gpurun_out/strip.log
timeout 60 ./bin/gol 6 16384 1024 128 0 --json --k 6 2>&1 | tail -1 >> gpurun_out/strip.log
timeout 60 ./bin/gol 6 16384 1024 128 0 --json --k 8 2>&1 | tail -1 >> gpurun_out/strip.log
cat gpurun_out/strip.log
