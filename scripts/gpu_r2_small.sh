mkdir -p gpurun_out; : > gpurun_out/small.log
run() { echo "# $*" >> gpurun_out/small.log; timeout 120 env "$@" 2>&1 | tail -1 | cut -c1-260 >> gpurun_out/small.log; }
run GOL_FLOW_FLAGS=0 ./bin/gol 6 4096 4096 256 0 --json --k 4
for t in 32 64 128 256; do for st in 11 16 21 28 42; do
run GOL_FLOW_FLAGS=4 ./bin/gol 6 4096 4096 $t 0 --json --k 4 --mode persistent --strip $st
done; done
for t in 64 128; do for st in 16 21 28; do
run GOL_FLOW_FLAGS=0 ./bin/gol 6 4096 4096 $t 0 --json --k 4 --mode persistent --strip $st
done; done
run GOL_FLOW_FLAGS=4 ./bin/gol 6 4096 4096 128 0 --json --k 2 --mode persistent --strip 21
run GOL_FLOW_FLAGS=4 ./bin/gol 6 4096 4096 128 0 --json --k 8 --mode persistent --strip 21
run GOL_FLOW_FLAGS=4 ./bin/gol 6 4096 4096 128 0 --json --k 8 --mode persistent --strip 42
grep -v "^#" gpurun_out/small.log | awk -F'"' '{print $0}' | cut -c1-200
