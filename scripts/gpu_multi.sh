set -x
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 1200 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi.log
tail -30 gpurun_out/pytest_multi.log
N=$(nvidia-smi -L | wc -l)
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_ours_$N.log 2>&1; echo "bench exit $?"
tail -1 gpurun_out/bench_ours_$N.log
: > gpurun_out/latency_$N.log
for cfg in "--k 4 --depth 4" "--k 4 --depth 16" "--k 4 --depth 32" "--k 8 --depth 32" "--mode persistent --k 4 --depth 4" "--mode persistent --k 4 --depth 16"; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 5 --warmup 3 --size 4096 --gens 4096 $cfg 2>&1 | tail -1 >> gpurun_out/latency_$N.log
done
cat gpurun_out/latency_$N.log | cut -c1-400
timeout 120 ./bin/gol 6 16384 1024 128 0 --json --ranks $N >> gpurun_out/single_proc_$N.log 2>&1
timeout 120 ./bin/gol 6 16384 1024 128 0 --json --ranks $N --mode persistent >> gpurun_out/single_proc_$N.log 2>&1
tail -4 gpurun_out/single_proc_$N.log
