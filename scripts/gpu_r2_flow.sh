# round 2, first GPU contact: full GPU test suite with the new interior loop + flow kernel, then
# fused (graph of per-pass launches) vs persistent (flow kernel) timings.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
: > gpurun_out/flow.log
run() { echo "# $*" >> gpurun_out/flow.log; timeout 120 "$@" 2>&1 | tail -1 >> gpurun_out/flow.log; }
run ./bin/gol 6 16384 1024 256 0 --json --k 4
run ./bin/gol 6 16384 1024 256 0 --json --k 4
run ./bin/gol 6 16384 1024 256 0 --json --k 4 --mode persistent
run ./bin/gol 6 16384 1024 256 0 --json --k 4 --mode persistent
for st in 32 48 56 60 63 64 72 96 128; do run ./bin/gol 6 16384 1024 256 0 --json --k 4 --mode persistent --strip $st; done
run ./bin/gol 6 16384 1024 256 0 --json --k 3 --mode persistent
run ./bin/gol 6 16384 1024 256 0 --json --k 2 --mode persistent
run ./bin/gol 6 16384 1024 128 0 --json --k 6 --mode persistent
run ./bin/gol 6 16384 1024 128 0 --json --k 8 --mode persistent
run ./bin/gol 6 4096 4096 256 0 --json --k 4
run ./bin/gol 6 4096 4096 256 0 --json --k 4 --mode persistent
for st in 8 12 16 24 32; do run ./bin/gol 6 4096 4096 256 0 --json --k 4 --mode persistent --strip $st; done
run ./bin/gol 6 65536 64 256 0 --json --k 4 --rows 8192
run ./bin/gol 6 65536 64 256 0 --json --k 4 --rows 8192 --mode persistent
run ./bin/gol 6 65536 16 256 0 --json --k 4
run ./bin/gol 6 65536 16 256 0 --json --k 4 --mode persistent
cat gpurun_out/flow.log
