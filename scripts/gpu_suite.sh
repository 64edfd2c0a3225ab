mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
: > gpurun_out/quick.log
run() { echo "# $*" >> gpurun_out/quick.log; env "$@" 2>&1 | tail -1 >> gpurun_out/quick.log; }
run GOL_PDL=0 timeout 120 ./bin/gol 6 4096 4096 256 0 --json --k 8
run GOL_PDL=0 timeout 120 ./bin/gol 6 4096 4096 128 0 --json --k 8
run GOL_PDL=0 timeout 120 ./bin/gol 6 4096 4096 256 0 --json --k 6
run GOL_PDL=1 timeout 120 ./bin/gol 6 4096 4096 256 0 --json --k 4
run GOL_PDL=0 timeout 120 ./bin/gol 6 4096 4096 256 0 --json --k 4 --mode persistent
run GOL_PDL=0 timeout 120 ./bin/gol 6 4096 4096 128 0 --json --k 8 --mode persistent
run GOL_PDL=0 timeout 120 ./bin/gol 6 16384 1024 256 0 --json --k 4 --mode persistent
cat gpurun_out/quick.log
./bin/gol 6 512 24 128 1 --ranks 2 --devices 1 --check --outdir gpurun_out > gpurun_out/plain_san.log 2>&1 && \
timeout 300 compute-sanitizer --tool memcheck --error-exitcode 9 ./bin/gol 6 512 24 128 1 --ranks 2 --devices 1 --check --outdir gpurun_out > gpurun_out/sanitizer_memcheck.log 2>&1; echo "memcheck exit $?" >> gpurun_out/sanitizer_memcheck.log
tail -4 gpurun_out/sanitizer_memcheck.log
