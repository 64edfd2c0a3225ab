mkdir -p gpurun_out
: > gpurun_out/tma.log
run() { echo "# $*" >> gpurun_out/tma.log; env "$@" 2>&1 | tail -1 >> gpurun_out/tma.log; }
timeout 300 python -m pytest tests -m gpu -x -q -k "tma" 2>&1 | tail -1 >> gpurun_out/tma.log
GOL_TMA_RANK=2 timeout 300 python -m pytest tests -m gpu -x -q -k "tma" 2>&1 | tail -1 >> gpurun_out/tma.log
run GOL_TMA_RANK=3 timeout 120 ./bin/gol 6 16384 1024 128 0 --json --k 4 --mode tma
run GOL_TMA_RANK=2 timeout 120 ./bin/gol 6 16384 1024 128 0 --json --k 4 --mode tma
run GOL_TMA_RANK=3 timeout 120 ./bin/gol 6 16384 1024 128 0 --json --k 2 --mode tma
run GOL_TMA_RANK=3 timeout 120 ./bin/gol 6 65536 256 128 0 --json --k 4 --mode tma
cat gpurun_out/tma.log
