mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "process_per_gpu" > gpurun_out/pytest_multi.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi.log
tail -3 gpurun_out/pytest_multi.log
OUT=gpurun_out/multi_$N.log; : > $OUT
tr() { local n=$1; shift; echo "# n=$n $*" >> $OUT
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $n "$@" 2>gpurun_out/last_err.log | tail -1 >> $OUT || { echo "FAILED" >> $OUT; tail -5 gpurun_out/last_err.log >> $OUT; }; }
tr $N --steps 5 --warmup 3 --size 4096 --gens 4096
tr $N --steps 5 --warmup 3 --size 4096 --gens 4096 --depth 16
tr $N --steps 10 --warmup 3
python - <<'PY'
import json
for l in open('gpurun_out/multi_2.log'):
    if l.startswith('{'):
        d=json.loads(l); print(round(d['value'],1), d['ms_per_step'], d['config'].get('strip_rows'), d['config'].get('exchange_depth'))
    else: print(l.strip())
PY
