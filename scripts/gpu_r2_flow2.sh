mkdir -p gpurun_out
: > gpurun_out/flow2.log
run() { echo "# $*" >> gpurun_out/flow2.log; timeout 120 env "$@" 2>&1 | tail -1 >> gpurun_out/flow2.log; }
B="./bin/gol 6 16384 1024 256 0 --json --k 4 --mode persistent"
run GOL_FLOW_FLAGS=0 $B
run GOL_FLOW_FLAGS=1 $B
run GOL_FLOW_FLAGS=2 $B
run GOL_FLOW_FLAGS=3 $B
run GOL_FLOW_FLAGS=0 $B --strip 56
run GOL_FLOW_FLAGS=1 $B --strip 56
run GOL_FLOW_FLAGS=2 $B --strip 56
run GOL_FLOW_FLAGS=0 GOL_FLOW_CTAS=148 $B
run GOL_FLOW_FLAGS=0 GOL_FLOW_CTAS=148 $B --strip 126
run GOL_FLOW_FLAGS=1 GOL_FLOW_CTAS=148 $B --strip 126
run GOL_FLOW_FLAGS=0 GOL_FLOW_CTAS=222 $B --strip 84
run GOL_FLOW_FLAGS=0 $B --strip 31
run GOL_FLOW_FLAGS=1 $B --strip 31
run GOL_FLOW_FLAGS=0 ./bin/gol 6 16384 1024 128 0 --json --k 4 --mode persistent
run GOL_FLOW_FLAGS=0 ./bin/gol 6 16384 1024 64 0 --json --k 4 --mode persistent
run GOL_FLOW_FLAGS=0 ./bin/gol 6 16384 1024 32 0 --json --k 4 --mode persistent
cat gpurun_out/flow2.log
