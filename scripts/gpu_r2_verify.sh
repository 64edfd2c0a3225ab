# what the driver runs at round end, plus one attempt at counters for the resident kernel
mkdir -p gpurun_out /tmp/ncu
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_final.log 2>&1; echo "bench exit $?"
tail -1 gpurun_out/bench_final.log | cut -c1-330
./bin/gol 6 4096 64 256 0 --mode resident > gpurun_out/plain_r.log 2>&1 && \
timeout 300 ncu --replay-mode application --section SpeedOfLight --section LaunchStats --section WarpStateStats --clock-control none -k regex:life_resident_kernel -c 1 -f -o /tmp/ncu/prof_r2_resident_app ./bin/gol 6 4096 64 256 0 --mode resident > gpurun_out/ncu_r_app.log 2>&1
tail -2 gpurun_out/ncu_r_app.log
ncu -i /tmp/ncu/prof_r2_resident_app.ncu-rep --page raw --csv 2>/dev/null | python3 -c "
import csv,sys
rows=list(csv.reader(sys.stdin))
if len(rows)>2:
    h=rows[0]; v=rows[2]
    for k in ('gpu__time_duration.sum','sm__throughput.avg.pct_of_peak_sustained_elapsed','smsp__inst_executed.sum','sm__warps_active.avg.pct_of_peak_sustained_active','launch__cluster_size','launch__grid_size','launch__block_size','launch__registers_per_thread','launch__shared_mem_per_block_dynamic','sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active','smsp__issue_active.avg.pct_of_peak_sustained_active'):
        if k in h: print(k, v[h.index(k)], rows[1][h.index(k)])
" > gpurun_out/ncu_r2_resident_app.txt 2>&1
cat gpurun_out/ncu_r2_resident_app.txt
