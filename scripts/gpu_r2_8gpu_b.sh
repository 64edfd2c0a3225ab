# short second 8-GPU session: the resident kernel on real NVLink rings + the latency config
mkdir -p gpurun_out
./bin/microbench_sync > gpurun_out/microbench_sync.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi_8b.log 2>&1; echo "pytest multi exit $?" >> gpurun_out/pytest_multi_8b.log
tail -3 gpurun_out/pytest_multi_8b.log
: > gpurun_out/multi8b.log
run() { echo "# $*" >> gpurun_out/multi8b.log; timeout 300 env "$@" 2>&1 | tail -1 >> gpurun_out/multi8b.log; }
T8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 8"
run GOL_X=0 $T8 --steps 10 --warmup 3 --size 4096 --gens 4096 --no-small
run GOL_X=0 $T8 --steps 5 --warmup 3 --size 8192 --gens 1024 --no-small
run GOL_X=0 ./bin/gol 6 4096 4096 256 0 --json --spawn 8
python3 - <<'PY'
import json
for line in open('gpurun_out/multi8b.log'):
    if line.startswith('#'): print(line.strip()[:150])
    if line.startswith('{'):
        d=json.loads(line)
        if 'metric' in d: print("  ", d['n_gpus'], d['config']['cells_per_gpu'], d['config'].get('mode'), 'GCUPS', round(d['value']), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), 'halo', d['phase']['halo_wait_us_per_gen'], d['verified'], 'clk', d['clocks'].get('sm_mhz'), d['clocks'].get('samples'))
        else: print("  ", line[:220])
PY
cat gpurun_out/microbench_sync.txt
