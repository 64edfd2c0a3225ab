mkdir -p gpurun_out; : > gpurun_out/strip2.log
run() { echo "# $*" >> gpurun_out/strip2.log; timeout 120 env "$@" 2>&1 | tail -1 | cut -c1-140 >> gpurun_out/strip2.log; }
for st in 0 496 500 503 504 505 509 512 520 336 252; do run GOL_X=1 ./bin/gol 6 65536 256 256 0 --json --k 4 --rows 32768 --strip $st; done
for st in 0 120 125 126 127 128 131 63 84; do run GOL_X=1 ./bin/gol 6 65536 256 256 0 --json --k 4 --rows 8192 --strip $st; done
for st in 0 61 62 63 64 65; do run GOL_X=1 ./bin/gol 6 16384 1024 256 0 --json --k 4 --strip $st; done
cat gpurun_out/strip2.log
