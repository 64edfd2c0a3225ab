# final single-GPU session of round 2: microbenchmarks, full test suite, bench lines, ncu captures
mkdir -p gpurun_out
./bin/microbench_sync > gpurun_out/microbench_sync.txt 2>&1; cat gpurun_out/microbench_sync.txt
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
: > gpurun_out/final1.log
run() { echo "# $*" >> gpurun_out/final1.log; timeout 600 env "$@" 2>&1 | tail -1 >> gpurun_out/final1.log; }
run GOL_X=0 python bench.py --steps 20 --warmup 5
run GOL_X=0 python bench.py --steps 5 --warmup 3 --size 4096 --gens 4096 --no-small
run GOL_X=0 python bench.py --steps 5 --warmup 3 --size 4096 --gens 4096 --no-small --mode fused
run GOL_X=0 python bench.py --steps 5 --warmup 3 --size 4096 --gens 4096 --no-small --mode persistent
run GOL_X=0 python bench.py --steps 5 --warmup 3 --size 8192 --gens 1024 --no-small --mode resident
run GOL_X=0 python bench.py --steps 5 --warmup 3 --size 8192 --gens 1024 --no-small --mode fused
run GOL_X=0 python bench.py --steps 5 --warmup 3 --no-small --mode persistent
run GOL_X=0 python bench.py --steps 5 --warmup 3 --no-small --mode tma
run GOL_REF_BUDGET_S=60 python bench.py --impl reference --steps 5 --warmup 3
python3 - <<'PY'
import json
for line in open('gpurun_out/final1.log'):
    if line.startswith('#'): print(line.strip()[:120])
    if line.startswith('{'):
        d=json.loads(line)
        if 'e2e' in d and 'config' in d: print("  ", d.get('impl'), d['config']['cells_per_gpu'], d['config'].get('mode'), 'GCUPS', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), 'steps', d['steps'], 'clk', d.get('clocks',{}).get('sm_mhz'))
        else: print("  ", line[:200])
PY
# ---- ncu: each capture right after the same command has exited 0 without ncu
N="ncu --set full --clock-control none --import-source on"
./bin/gol 6 16384 64 256 0 --k 4 --mode fused > gpurun_out/plain_f.log 2>&1 && \
$N -k regex:life_fused_kernel -s 6 -c 1 -f -o gpurun_out/prof_r2_fused_k4 ./bin/gol 6 16384 64 256 0 --k 4 --mode fused > gpurun_out/ncu_f.log 2>&1
tail -1 gpurun_out/ncu_f.log
./bin/gol 6 4096 256 256 0 --mode resident > gpurun_out/plain_r.log 2>&1 && \
$N -k regex:life_resident_kernel -c 1 -f -o gpurun_out/prof_r2_resident ./bin/gol 6 4096 256 256 0 --mode resident > gpurun_out/ncu_r.log 2>&1
tail -1 gpurun_out/ncu_r.log
./bin/gol 6 16384 128 256 0 --k 4 --mode persistent > gpurun_out/plain_p.log 2>&1 && \
$N -k regex:life_flow_kernel -c 1 -f -o gpurun_out/prof_r2_flow_k4 ./bin/gol 6 16384 128 256 0 --k 4 --mode persistent > gpurun_out/ncu_p.log 2>&1
tail -1 gpurun_out/ncu_p.log
python scripts/ncu_secondary.py > gpurun_out/plain_s.log 2>&1 && \
ncu --set full --clock-control none -k "regex:pack_kernel|unpack_kernel|population_partial|population_final|step_rows|format_rows|life_generic|halo_prime|pattern_init" -c 24 -f -o gpurun_out/prof_r2_secondary python scripts/ncu_secondary.py > gpurun_out/ncu_s.log 2>&1
tail -1 gpurun_out/ncu_s.log
ls -la gpurun_out/prof_r2_*.ncu-rep
