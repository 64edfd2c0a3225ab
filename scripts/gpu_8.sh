# 8-GPU session: correctness on all GPUs, weak scaling 1/2/4/8, strong scaling 65536^2,
# 131072^2 persistent run, 4096^2 latency config, NCCL baseline, reference (steelman + as-shipped)
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
OUT=gpurun_out/multi_$N.log
: > $OUT
tr() { # tr <nproc> <args...>
  local n=$1; shift
  echo "# n=$n $*" >> $OUT
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n "$@" 2>gpurun_out/last_err.log | tail -1 >> $OUT || { echo "FAILED" >> $OUT; tail -5 gpurun_out/last_err.log >> $OUT; }
}
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi_$N.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi_$N.log
tail -3 gpurun_out/pytest_multi_$N.log
# weak scaling, headline config
for n in 1 2 4 8; do [ $n -le $N ] && tr $n --steps 10 --warmup 3; done
# strong scaling: 65536^2 global
for n in 1 2 4 8; do [ $n -le $N ] && tr $n --steps 3 --warmup 3 --size 65536 --rows $((65536 / n)) --gens 256; done
# 131072^2 global on all GPUs: persistent kernel and graph of fused passes
tr $N --steps 3 --warmup 3 --size 131072 --rows $((131072 / N)) --gens 128 --mode persistent
tr $N --steps 3 --warmup 3 --size 131072 --rows $((131072 / N)) --gens 128
tr $N --steps 3 --warmup 3 --size 131072 --rows $((131072 / N)) --gens 128 --k 8
# 4096^2 per GPU, latency bound
for n in 1 2 4 8; do [ $n -le $N ] && tr $n --steps 5 --warmup 3 --size 4096 --gens 4096; done
tr $N --steps 5 --warmup 3 --size 4096 --gens 4096 --depth 16
tr $N --steps 5 --warmup 3 --size 4096 --gens 4096 --mode persistent
tr $N --steps 5 --warmup 3 --size 4096 --gens 4096 --mode persistent --depth 16
# NCCL-halo baseline (library collectives + unfused stencil)
tr $N --impl nccl --steps 3 --warmup 3
tr 1 --impl nccl --steps 3 --warmup 3
# single process driving all GPUs (peer access)
echo "# single process, $N ranks" >> $OUT
timeout 120 ./bin/gol 6 16384 1024 256 0 --json --ranks $N 2>&1 | tail -1 >> $OUT
# reference under its own launcher on all GPUs
echo "# reference n=$N" >> $OUT
GOL_REF_BUDGET_S=20 timeout 600 python bench.py --impl reference --gpus $N --steps 5 --warmup 3 2>&1 | tail -1 >> $OUT
cut -c1-330 $OUT
