mkdir -p gpurun_out; : > gpurun_out/b3.log
python bench.py --steps 3 --warmup 3 --size 65536 --rows 32768 --no-small >> gpurun_out/b3.log 2>&1
python bench.py --steps 10 --warmup 3 >> gpurun_out/b3.log 2>&1
python bench.py --steps 3 --warmup 3 --scaling strong >> gpurun_out/b3.log 2>&1
python3 - <<'PY'
import json
for line in open('gpurun_out/b3.log'):
    if line.startswith('{') and '"metric"' in line:
        d=json.loads(line)
        print("SUMMARY", d['config']['cells_per_gpu'], d['scaling'], round(d['value']), d['ms_per_step'], 'e2e', round(d['e2e']['value']), d['e2e']['ms_per_step'], d['e2e']['serial_ms_per_step'])
PY
