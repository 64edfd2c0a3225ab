mkdir -p gpurun_out; : > gpurun_out/multi2c.log
run() { echo "# $*" >> gpurun_out/multi2c.log; timeout 300 env "$@" 2>&1 | tail -2 >> gpurun_out/multi2c.log; }
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
run GOL_X=0 $TR -m gol_b200.cli 6 65536 256 256 0 --rows 32768
run GOL_X=0 $TR -m gol_b200.cli 6 65536 1024 256 0 --rows 32768
run GOL_X=0 ./bin/gol 6 65536 1024 256 0 --json --k 4 --rows 32768 --ranks 2
run GOL_X=0 ./bin/gol 6 65536 1024 256 0 --json --k 4 --rows 32768 --ranks 1
run GOL_X=0 python bench.py --steps 3 --warmup 3 --size 65536 --rows 32768 --no-small
cut -c1-330 gpurun_out/multi2c.log
