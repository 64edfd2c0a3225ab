set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_ours.log 2>&1; echo "bench exit $?"
tail -2 gpurun_out/bench_ours.log
: > gpurun_out/ksweep.log
for k in 1 2 3 4 6 8; do timeout 120 ./bin/gol 6 16384 1024 128 0 --json --k $k 2>&1 | tail -1 >> gpurun_out/ksweep.log; done
for t in 64 256; do timeout 120 ./bin/gol 6 16384 1024 $t 0 --json --k 4 2>&1 | tail -1 >> gpurun_out/ksweep.log; done
timeout 120 ./bin/gol 6 16384 1024 128 0 --json --mode persistent --k 4 2>&1 | tail -1 >> gpurun_out/ksweep.log
timeout 120 ./bin/gol 6 16384 1024 128 0 --json --k 4 --depth 16 2>&1 | tail -1 >> gpurun_out/ksweep.log
timeout 120 ./bin/gol 6 65536 256 128 0 --json --k 4 --rows 32768 2>&1 | tail -1 >> gpurun_out/ksweep.log
cat gpurun_out/ksweep.log
for n in 512 1024 2048; do ( time timeout 200 baseline/_ref/bin/mpirun -n 1 baseline/_ref/bin/gol_ref 1 $n 1 512 0 | tail -c 200 ) > gpurun_out/ref_asshipped_$n.log 2>&1; tail -5 gpurun_out/ref_asshipped_$n.log; done
./bin/gol 6 16384 64 128 0 --k 4 > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:life_fused -s 6 -c 2 -o gpurun_out/prof_fused_k4 ./bin/gol 6 16384 64 128 0 --k 4 > gpurun_out/ncu1.log 2>&1
tail -3 gpurun_out/ncu1.log
