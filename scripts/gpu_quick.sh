mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
: > gpurun_out/quick.log
run() { echo "# $*" >> gpurun_out/quick.log; env "$@" 2>&1 | tail -1 >> gpurun_out/quick.log; }
run GOL_PDL=0 timeout 120 ./bin/gol 6 16384 1024 256 0 --json --k 4
run GOL_PDL=0 timeout 120 ./bin/gol 6 4096 4096 256 0 --json --k 4
run GOL_PDL=0 timeout 120 ./bin/gol 6 4096 4096 256 0 --json --k 4 --strip 12
run GOL_PDL=0 timeout 120 ./bin/gol 6 4096 4096 256 0 --json --k 4 --mode persistent --strip 12
run GOL_PDL=0 timeout 120 ./bin/gol 6 4096 4096 256 0 --json --k 4 --ranks 2 --devices 1
cat gpurun_out/quick.log
