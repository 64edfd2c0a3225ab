mkdir -p gpurun_out
: > gpurun_out/occ.log
run() { echo "# $*" >> gpurun_out/occ.log; timeout 120 env "$@" 2>&1 | tail -1 | cut -c1-200 >> gpurun_out/occ.log; }
for mb in 0 3 4; do
  run GOL_EXP_MINB=$mb ./bin/gol 6 16384 1024 256 0 --json --k 2
  run GOL_EXP_MINB=$mb ./bin/gol 6 16384 1024 256 0 --json --k 3
done
for st in 40 48 56 72; do run GOL_EXP_MINB=3 ./bin/gol 6 16384 1024 256 0 --json --k 2 --strip $st; done
for st in 40 48 56 72; do run GOL_EXP_MINB=3 ./bin/gol 6 16384 1024 256 0 --json --k 3 --strip $st; done
cat gpurun_out/occ.log
