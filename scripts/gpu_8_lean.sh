mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
OUT=gpurun_out/multi_${N}_final.log; : > $OUT
tr() { local n=$1; shift; echo "# n=$n $*" >> $OUT
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n "$@" 2>gpurun_out/last_err.log | tail -1 >> $OUT || { echo "FAILED" >> $OUT; tail -5 gpurun_out/last_err.log >> $OUT; }; }
tr $N --steps 10 --warmup 3
tr $N --steps 5 --warmup 3 --size 4096 --gens 4096
tr $N --steps 3 --warmup 3 --size 65536 --rows $((65536 / N)) --gens 256
python - <<'PY'
import json,glob
for f in glob.glob('gpurun_out/multi_*_final.log'):
    for l in open(f):
        if l.startswith('{'):
            d=json.loads(l); print(round(d['value'],1), d['n_gpus'], d['ms_per_step'], d['config'].get('cells_per_gpu'), 'e2e', round(d['e2e']['value'],1), d['clocks'])
        else: print(l.strip())
PY
