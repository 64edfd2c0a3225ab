mkdir -p gpurun_out; : > gpurun_out/multi2b.log
run() { echo "# $*" >> gpurun_out/multi2b.log; timeout 300 env "$@" 2>&1 | tail -1 >> gpurun_out/multi2b.log; }
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2"
run GOL_X=0 $T --steps 5 --warmup 3 --scaling strong --no-small
run GOL_X=0 ./bin/gol 6 65536 256 256 0 --json --k 4 --rows 32768 --ranks 2
run GOL_X=0 ./bin/gol 6 65536 256 256 0 --json --k 4 --rows 32768 --ranks 1
run GOL_X=0 $T --steps 5 --warmup 3 --no-small
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -2 >> gpurun_out/multi2b.log
python3 - <<'PY' >> gpurun_out/multi2b.log
import json
for line in open('gpurun_out/multi2b.log'):
    if line.startswith('{') and '"metric"' in line:
        d=json.loads(line)
        print("SUMMARY", d['config']['cells_per_gpu'], d['scaling'], round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']), d['e2e']['ms_per_step'])
PY
grep -v '"metric"' gpurun_out/multi2b.log | cut -c1-200
