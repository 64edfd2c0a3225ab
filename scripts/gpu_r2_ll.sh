mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
GOL_FORCE_LL=1 timeout 600 python -m pytest tests/test_gpu_property.py tests/test_gpu_wide_oracle.py -m gpu -x -q > gpurun_out/pytest_ll.log 2>&1; echo "pytest LL exit $?" >> gpurun_out/pytest_ll.log
tail -3 gpurun_out/pytest_ll.log
: > gpurun_out/ll.log
for f in 0 1; do
GOL_FORCE_LL=$f ./bin/gol 6 16384 1024 256 0 --json --k 4 | tail -1 | cut -c1-220 >> gpurun_out/ll.log
GOL_FORCE_LL=$f ./bin/gol 6 4096 4096 256 0 --json --k 4 | tail -1 | cut -c1-220 >> gpurun_out/ll.log
GOL_FORCE_LL=$f ./bin/gol 6 4096 4096 256 0 --json --k 4 --ranks 2 --devices 1 | tail -1 | cut -c1-220 >> gpurun_out/ll.log
done
cat gpurun_out/ll.log
