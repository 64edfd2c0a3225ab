set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.log 2>&1; tail -3 gpurun_out/pytest_noalloc_alone.log
GOL_FUSED_VARIANT=1 timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "fused or persistent or larger" > gpurun_out/pytest_v1.log 2>&1; tail -3 gpurun_out/pytest_v1.log
: > gpurun_out/ksweep.log
for v in 0 1; do for t in 128 256; do GOL_FUSED_VARIANT=$v timeout 120 ./bin/gol 6 16384 1024 $t 0 --json --k 4 2>&1 | tail -1 >> gpurun_out/ksweep.log; done; done
for k in 2 3 6 8; do timeout 120 ./bin/gol 6 16384 1024 128 0 --json --k $k 2>&1 | tail -1 >> gpurun_out/ksweep.log; done
timeout 120 ./bin/gol 6 16384 1024 128 0 --json --mode persistent --k 4 2>&1 | tail -1 >> gpurun_out/ksweep.log
for v in 0 1; do GOL_FUSED_VARIANT=$v timeout 120 ./bin/gol 6 65536 256 128 0 --json --k 4 --rows 32768 2>&1 | tail -1 >> gpurun_out/ksweep.log; done
cat gpurun_out/ksweep.log
GOL_FUSED_VARIANT=1 ./bin/gol 6 16384 64 128 0 --k 4 > gpurun_out/plain.log 2>&1 && \
GOL_FUSED_VARIANT=1 ncu --set full --clock-control none --import-source on -k regex:life_fused -s 6 -c 2 -o gpurun_out/prof_fused_k4_v1 ./bin/gol 6 16384 64 128 0 --k 4 > gpurun_out/ncu1.log 2>&1
tail -2 gpurun_out/ncu1.log
