set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
: > gpurun_out/ksweep.log
for v in 0 2; do GOL_FUSED_VARIANT=$v timeout 120 ./bin/gol 6 16384 1024 128 0 --json --k 4 2>&1 | tail -1 >> gpurun_out/ksweep.log; done
GOL_FUSED_VARIANT=3 timeout 120 ./bin/gol 6 16384 1024 128 0 --json --k 3 2>&1 | tail -1 >> gpurun_out/ksweep.log
timeout 120 ./bin/gol 6 16384 1024 256 0 --json --k 4 2>&1 | tail -1 >> gpurun_out/ksweep.log
timeout 120 ./bin/gol 6 16384 1024 0 0 --json --mode persistent --k 4 2>&1 | tail -1 >> gpurun_out/ksweep.log
timeout 120 ./bin/gol 6 16384 1024 128 0 --json --mode persistent --k 4 2>&1 | tail -1 >> gpurun_out/ksweep.log
timeout 120 ./bin/gol 6 4096 4096 128 0 --json --k 4 2>&1 | tail -1 >> gpurun_out/ksweep.log
timeout 120 ./bin/gol 6 4096 4096 0 0 --json --k 4 --mode persistent 2>&1 | tail -1 >> gpurun_out/ksweep.log
timeout 120 ./bin/gol 6 65536 256 128 0 --json --k 4 2>&1 | tail -1 >> gpurun_out/ksweep.log
cat gpurun_out/ksweep.log
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_ours.log 2>&1; echo "bench exit $?"
tail -1 gpurun_out/bench_ours.log | cut -c1-300
timeout 900 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref.log 2>&1; echo "ref exit $?"
tail -1 gpurun_out/bench_ref.log
