mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_property.py -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
: > gpurun_out/flow4.log
run() { echo "# $*" >> gpurun_out/flow4.log; timeout 120 env "$@" 2>&1 | tail -1 >> gpurun_out/flow4.log; }
B="./bin/gol 6 16384 1024 256 0 --json --k 4 --mode persistent"
run GOL_FLOW_FLAGS=0 ./bin/gol 6 16384 1024 256 0 --json --k 4
run GOL_FLOW_FLAGS=0 $B
run GOL_FLOW_FLAGS=1 $B
run GOL_FLOW_FLAGS=4 $B
run GOL_FLOW_FLAGS=5 $B
for st in 24 31 40 48 56; do run GOL_FLOW_FLAGS=4 $B --strip $st; done
for st in 31 48 56 60 64; do run GOL_FLOW_FLAGS=0 $B --strip $st; done
run GOL_FLOW_FLAGS=0 ./bin/gol 6 4096 4096 256 0 --json --k 4
run GOL_FLOW_FLAGS=0 ./bin/gol 6 4096 4096 256 0 --json --k 4 --mode persistent
run GOL_FLOW_FLAGS=4 ./bin/gol 6 4096 4096 256 0 --json --k 4 --mode persistent
run GOL_FLOW_FLAGS=4 ./bin/gol 6 4096 4096 256 0 --json --k 4 --mode persistent --strip 8
run GOL_FLOW_FLAGS=4 ./bin/gol 6 4096 4096 256 0 --json --k 4 --mode persistent --strip 24
run GOL_FLOW_FLAGS=0 ./bin/gol 6 65536 16 256 0 --json --k 4 --mode persistent
run GOL_FLOW_FLAGS=4 ./bin/gol 6 65536 16 256 0 --json --k 4 --mode persistent --strip 256
cat gpurun_out/flow4.log
