mkdir -p gpurun_out /tmp/ncu
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
: > gpurun_out/final3.log
run() { echo "# $*" >> gpurun_out/final3.log; timeout 600 env "$@" 2>&1 | tail -1 >> gpurun_out/final3.log; }
run GOL_X=0 python bench.py --steps 5 --warmup 3 --size 4096 --gens 4096 --no-small
run GOL_X=0 python bench.py --steps 5 --warmup 3 --size 8192 --gens 1024 --no-small --mode resident
run GOL_X=0 python bench.py --steps 5 --warmup 3 --no-small --mode tma
run GOL_X=0 ./bin/gol 6 4096 4096 256 0 --json
run GOL_X=0 ./bin/gol 6 8192 1024 256 0 --json --mode resident
run GOL_X=0 ./bin/gol 6 8192 1024 256 0 --json --mode fused
run GOL_X=0 ./bin/gol 6 2048 8192 256 0 --json
run GOL_X=0 ./bin/gol 6 2048 8192 256 0 --json --mode fused
python3 - <<'PY'
import json
for line in open('gpurun_out/final3.log'):
    if line.startswith('#'): print(line.strip()[:120])
    if line.startswith('{'):
        d=json.loads(line)
        if 'e2e' in d and 'config' in d: print("  ", d.get('impl'), d['config']['cells_per_gpu'], d['config'].get('mode'), 'GCUPS', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'steps', d['steps'])
        else: print("  ", line[:160])
    elif not line.startswith('#'): print("  ?", line[:200].strip())
PY
N="ncu --set full --clock-control none --import-source on"
./bin/gol 6 16384 64 256 0 --k 4 --mode fused > gpurun_out/plain_f.log 2>&1 && \
$N -k regex:life_fused_kernel -s 6 -c 1 -f -o /tmp/ncu/prof_r2_fused_k4 ./bin/gol 6 16384 64 256 0 --k 4 --mode fused > gpurun_out/ncu_f.log 2>&1
./bin/gol 6 4096 256 256 0 --mode resident > gpurun_out/plain_r.log 2>&1 && \
$N -k regex:life_resident_kernel -c 1 -f -o /tmp/ncu/prof_r2_resident ./bin/gol 6 4096 256 256 0 --mode resident > gpurun_out/ncu_r.log 2>&1
./bin/gol 6 16384 128 256 0 --k 4 --mode persistent > gpurun_out/plain_p.log 2>&1 && \
$N -k regex:life_flow_kernel -c 1 -f -o /tmp/ncu/prof_r2_flow_k4 ./bin/gol 6 16384 128 256 0 --k 4 --mode persistent > gpurun_out/ncu_p.log 2>&1
python scripts/ncu_secondary.py > gpurun_out/plain_s.log 2>&1 && \
ncu --set full --clock-control none -k "regex:pack_kernel|unpack_kernel|population_partial|population_final|step_rows|format_rows|life_generic|halo_prime|pattern_init" -c 24 -f -o /tmp/ncu/prof_r2_secondary python scripts/ncu_secondary.py > gpurun_out/ncu_s.log 2>&1
ls -la /tmp/ncu
python scripts/ncu_summary.py /tmp/ncu/prof_r2_fused_k4.ncu-rep "life_fused_kernel<K=4> - one pass (4 generations) of a 16384x16384 slab (round 2 kernel: pointer-increment interior loop)" > gpurun_out/ncu_r2_fused_k4.md
python scripts/ncu_summary.py /tmp/ncu/prof_r2_resident.ncu-rep "life_resident_kernel - 256 generations of a 4096x4096 slab in one launch (strips in shared memory, clusters of 2)" > gpurun_out/ncu_r2_resident.md
python scripts/ncu_summary.py /tmp/ncu/prof_r2_flow_k4.ncu-rep "life_flow_kernel<K=4> - 128 generations of a 16384x16384 slab in one launch (tickets + per-strip progress counters)" > gpurun_out/ncu_r2_flow_k4.md
python scripts/ncu_summary.py /tmp/ncu/prof_r2_secondary.ncu-rep "Secondary kernels (scripts/ncu_secondary.py)" --all > gpurun_out/ncu_r2_secondary.md
cp /tmp/ncu/prof_r2_fused_k4.ncu-rep gpurun_out/ 2>/dev/null
head -40 gpurun_out/ncu_r2_fused_k4.md
