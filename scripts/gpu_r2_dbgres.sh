mkdir -p gpurun_out
GOL_B200_LIB=debug timeout 600 python -m pytest -x -q -m gpu -p no:cacheprovider tests/test_gpu_wide_oracle.py tests/test_gpu_kernels.py -k "resident" > gpurun_out/pytest_dbgres.log 2>&1; echo "exit $?" >> gpurun_out/pytest_dbgres.log
tail -4 gpurun_out/pytest_dbgres.log
timeout 600 python -m pytest -x -q -m gpu tests/test_gpu_wide_oracle.py tests/test_gpu_kernels.py tests/test_gpu_debug_bounds.py > gpurun_out/pytest_quick.log 2>&1; echo "exit $?" >> gpurun_out/pytest_quick.log
tail -3 gpurun_out/pytest_quick.log
for i in 1 2 3; do ./bin/gol 6 4096 4096 256 0 --json --check 2>&1 | tail -2 | cut -c1-120; done
