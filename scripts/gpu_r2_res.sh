mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_wide_oracle.py -m gpu -x -q -k resident > gpurun_out/pytest_res.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_res.log
tail -3 gpurun_out/pytest_res.log
: > gpurun_out/res.log
run() { echo "# $*" >> gpurun_out/res.log; timeout 60 env "$@" 2>&1 | tail -1 | cut -c1-230 >> gpurun_out/res.log; }
for t in 128 256 512; do for k in 2 4 8; do
run GOL_RESIDENT_CLUSTER=1 ./bin/gol 6 4096 4096 $t 0 --json --mode resident --k $k
done; done
for cs in 2 4 8; do run GOL_RESIDENT_CLUSTER=$cs ./bin/gol 6 4096 4096 512 0 --json --mode resident --k 4; done
run GOL_X=0 ./bin/gol 6 8192 1024 512 0 --json --mode resident --k 4 --rows 8192
run GOL_X=0 ./bin/gol 6 2048 8192 512 0 --json --mode resident --k 4
cat gpurun_out/res.log
