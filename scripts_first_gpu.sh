set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_ours.log 2>&1; echo "bench exit $?"
tail -3 gpurun_out/bench_ours.log
for k in 1 2 4 8; do timeout 120 ./bin/gol 6 16384 256 128 0 --json --k $k 2>&1 | tail -1; done > gpurun_out/ksweep.log 2>&1
timeout 120 ./bin/gol 6 16384 256 128 0 --json --mode persistent --k 4 >> gpurun_out/ksweep.log 2>&1
timeout 120 ./bin/gol 6 16384 64 256 0 --json --mode generic >> gpurun_out/ksweep.log 2>&1
cat gpurun_out/ksweep.log
( time timeout 300 baseline/_ref/bin/mpirun -n 1 baseline/_ref/bin/gol_ref 1 16384 2 512 0 | tail -c 300 ) > gpurun_out/ref_asshipped.log 2>&1
( time timeout 300 baseline/_ref/bin/mpirun -n 1 baseline/_ref/bin/gol_ref_noprintf 1 16384 64 512 0 | tail -c 300 ) > gpurun_out/ref_noprintf.log 2>&1
cat gpurun_out/ref_asshipped.log gpurun_out/ref_noprintf.log
