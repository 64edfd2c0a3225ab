mkdir -p gpurun_out
: > gpurun_out/flow5.log
run() { echo "# $*" >> gpurun_out/flow5.log; timeout 120 env "$@" 2>&1 | tail -1 | cut -c1-330 >> gpurun_out/flow5.log; }
run GOL_FLOW_FLAGS=0 ./bin/gol 6 16384 1024 256 0 --json --k 4
B="./bin/gol 6 16384 1024 256 0 --json --k 4 --mode persistent"
for st in 36 40 44; do run GOL_FLOW_FLAGS=4 $B --strip $st; done
run GOL_FLOW_FLAGS=0 $B
# 8-GPU strong-scaling slab shape and the 131072-wide slab
S="./bin/gol 6 65536 256 256 0 --json --k 4 --rows 8192"
run GOL_FLOW_FLAGS=0 $S
for st in 64 96 128 192 256; do run GOL_FLOW_FLAGS=4 $S --mode persistent --strip $st; done
T="./bin/gol 6 131072 64 256 0 --json --k 4 --rows 16384"
run GOL_FLOW_FLAGS=0 $T
for st in 128 256 512; do run GOL_FLOW_FLAGS=4 $T --mode persistent --strip $st; done
U="./bin/gol 6 65536 64 256 0 --json --k 4"
run GOL_FLOW_FLAGS=0 $U
for st in 128 256 512; do run GOL_FLOW_FLAGS=4 $U --mode persistent --strip $st; done
cat gpurun_out/flow5.log
