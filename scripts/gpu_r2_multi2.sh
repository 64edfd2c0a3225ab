mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest multi exit $?" >> gpurun_out/pytest_multi.log
tail -5 gpurun_out/pytest_multi.log
: > gpurun_out/multi2.log
run() { echo "# $*" >> gpurun_out/multi2.log; timeout 300 env "$@" 2>&1 | tail -1 >> gpurun_out/multi2.log; }
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2"
run GOL_NO_LL=0 $T --steps 10 --warmup 3
run GOL_NO_LL=1 $T --steps 10 --warmup 3 --no-small
run GOL_NO_LL=0 $T --steps 5 --warmup 3 --size 4096 --gens 4096 --no-small
run GOL_NO_LL=1 $T --steps 5 --warmup 3 --size 4096 --gens 4096 --no-small
run GOL_NO_LL=0 $T --steps 5 --warmup 3 --size 4096 --gens 4096 --no-small --depth 16
run GOL_NO_LL=0 $T --steps 5 --warmup 3 --scaling strong --no-small
run GOL_NO_LL=0 ./bin/gol 6 16384 1024 256 0 --json --k 4 --ranks 2
run GOL_NO_LL=0 ./bin/gol 6 4096 4096 256 0 --json --k 4 --ranks 2
run GOL_NO_LL=1 ./bin/gol 6 4096 4096 256 0 --json --k 4 --ranks 2
cat gpurun_out/multi2.log | cut -c1-1500
