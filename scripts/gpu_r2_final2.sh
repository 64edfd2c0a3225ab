mkdir -p gpurun_out
GOL_B200_LIB=debug timeout 900 python -m pytest -x -q -m gpu -p no:cacheprovider tests/test_gpu_property.py tests/test_gpu_wide_oracle.py tests/test_gpu_kernels.py --deselect tests/test_gpu_kernels.py::test_slab_larger_than_2_pow_31_cells --deselect tests/test_gpu_kernels.py::test_native_cli_on_gpu > gpurun_out/pytest_debug.log 2>&1
tail -30 gpurun_out/pytest_debug.log | cut -c1-300
python bench.py --steps 3 --warmup 3 --size 8192 --gens 1024 --no-small --mode resident > gpurun_out/b8192.log 2>&1; tail -5 gpurun_out/b8192.log | cut -c1-400
: > gpurun_out/res3.log
run() { echo "# $*" >> gpurun_out/res3.log; timeout 60 env "$@" 2>&1 | tail -1 | cut -c1-150 >> gpurun_out/res3.log; }
for cs in 1 2 4 8; do run GOL_RESIDENT_CLUSTER=$cs ./bin/gol 6 4096 4096 256 0 --json --mode resident --k 4; done
run GOL_X=0 ./bin/gol 6 4096 4096 256 0 --json --mode resident --k 4
run GOL_X=0 ./bin/gol 6 4096 4096 256 0 --json
cat gpurun_out/res3.log
