set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "tma" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
: > gpurun_out/ksweep.log
run() { echo "# $*" >> gpurun_out/ksweep.log; env "$@" 2>&1 | tail -1 >> gpurun_out/ksweep.log; }
run GOL_PDL=0 timeout 120 ./bin/gol 6 16384 1024 256 0 --json --k 4
run GOL_PDL=0 timeout 120 ./bin/gol 6 16384 1024 128 0 --json --k 4 --mode tma
run GOL_PDL=1 timeout 120 ./bin/gol 6 16384 1024 128 0 --json --k 4 --mode tma
run GOL_PDL=0 timeout 120 ./bin/gol 6 16384 1024 128 0 --json --k 3 --mode tma
run GOL_PDL=0 timeout 120 ./bin/gol 6 16384 1024 128 0 --json --k 2 --mode tma
run GOL_PDL=0 timeout 120 ./bin/gol 6 65536 256 128 0 --json --k 4 --mode tma
run GOL_PDL=0 timeout 120 ./bin/gol 6 65536 256 256 0 --json --k 4
run GOL_PDL=0 timeout 120 ./bin/gol 6 16384 1024 128 0 --json --k 4 --mode tma --strip 63
run GOL_PDL=0 timeout 120 ./bin/gol 6 16384 1024 128 0 --json --k 4 --mode tma --strip 120
cat gpurun_out/ksweep.log
./bin/gol 6 16384 64 128 0 --k 4 --mode tma > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:life_fused_tma -s 6 -c 2 -o gpurun_out/prof_fused_tma_k4 ./bin/gol 6 16384 64 128 0 --k 4 --mode tma > gpurun_out/ncu1.log 2>&1
tail -2 gpurun_out/ncu1.log
