# the 8-GPU session of round 2: correctness on real NVLink peers, weak + strong scaling of both arms,
# native multi-rank drivers, latency and large configs in graph vs persistent mode
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_2d.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi_2d.log 2>&1; echo "pytest multi exit $?" >> gpurun_out/pytest_multi_2d.log
tail -3 gpurun_out/pytest_multi_2d.log
: > gpurun_out/multi2d.log
run() { echo "# $*" >> gpurun_out/multi2d.log; timeout 600 env "$@" 2>&1 | tail -1 >> gpurun_out/multi2d.log; }
T8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2"
T4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2"
run GOL_X=0 $T8 --steps 3 --warmup 3
run GOL_X=0 $T8 --steps 3 --warmup 3 --scaling strong --no-small
run GOL_X=0 $T4 --steps 3 --warmup 3 --scaling strong --no-small
run GOL_X=0 $T4 --steps 3 --warmup 3 --no-small
run GOL_X=0 $T8 --steps 5 --warmup 3 --size 4096 --gens 4096 --no-small
run GOL_X=0 $T8 --steps 5 --warmup 3 --size 4096 --gens 4096 --no-small --mode persistent
run GOL_X=0 $T8 --steps 5 --warmup 3 --size 131072 --rows 16384 --gens 256 --no-small
run GOL_X=0 $T8 --steps 5 --warmup 3 --size 131072 --rows 16384 --gens 256 --no-small --mode persistent
run GOL_X=0 $T8 --steps 5 --warmup 3 --no-small --mode persistent
run GOL_X=0 ./bin/gol 6 16384 1024 256 0 --json --ranks 2
run GOL_X=0 ./bin/gol 6 16384 1024 256 0 --json --spawn 2
run GOL_X=0 ./bin/gol 6 16384 1024 256 0 --json --ranks 2
run GOL_X=0 ./bin/gol 6 16384 1024 256 0 --json --spawn 2
# reference, strong scaling (rectangular variant) and weak, short runs for the table
for n in 2; do run GOL_REF_BUDGET_S=120 python bench.py --impl reference --gpus $n --scaling strong --gens 64 --steps 2; done
run GOL_REF_BUDGET_S=120 python bench.py --impl reference --gpus 2 --size 32768 --gens 64 --steps 2
run GOL_REF_BUDGET_S=120 python bench.py --impl reference --gpus 2 --gens 64 --steps 2
run GOL_X=0 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus 2 --impl nccl --steps 2
python3 - <<'PY'
import json
for line in open('gpurun_out/multi2d.log'):
    if line.startswith('#'): print(line.strip()[:150])
    if line.startswith('{'):
        try: d=json.loads(line)
        except Exception as e: print("BAD", line[:200]); continue
        if 'metric' in d and 'config' in d and 'e2e' in d:
            print("  ", d.get('impl'), d['n_gpus'], d['config']['cells_per_gpu'], d.get('scaling'), d['config'].get('mode'), 'GCUPS', round(d['value']), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']), d['e2e'].get('ms_per_step'), 'steps', d['steps'], 'clk', d.get('clocks',{}).get('sm_mhz'), d.get('clocks',{}).get('samples'), 'halo', d.get('phase',{}).get('halo_wait_us_per_gen'))
        else: print("  ", json.dumps(d)[:300])
PY
