mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest multi exit $?" >> gpurun_out/pytest_multi.log
tail -3 gpurun_out/pytest_multi.log
: > gpurun_out/res2.log
run() { echo "# $*" >> gpurun_out/res2.log; timeout 300 env "$@" 2>&1 | tail -1 >> gpurun_out/res2.log; }
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2"
run GOL_X=0 $T --steps 5 --warmup 3 --size 4096 --gens 4096 --no-small
run GOL_X=0 $T --steps 5 --warmup 3 --size 4096 --gens 4096 --no-small --k 8
run GOL_X=0 ./bin/gol 6 4096 4096 256 0 --json --ranks 2
run GOL_X=0 ./bin/gol 6 4096 4096 256 0 --json --spawn 2
python3 - <<'PY'
import json
for line in open('gpurun_out/res2.log'):
    if line.startswith('#'): print(line.strip()[:150])
    if line.startswith('{'):
        d=json.loads(line)
        if 'metric' in d: print("  ", d['n_gpus'], d['config']['cells_per_gpu'], d['config'].get('mode'), d['config'].get('k'), 'GCUPS', round(d['value']), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), 'halo', d['phase']['halo_wait_us_per_gen'], d['verified'])
        else: print("  ", line[:220])
PY
