mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_ours.log 2>&1; echo "bench exit $?"
tail -1 gpurun_out/bench_ours.log
timeout 900 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_ref.log 2>&1; echo "ref exit $?"
tail -1 gpurun_out/bench_ref.log
timeout 600 python bench.py --scaling strong --steps 5 --warmup 3 > gpurun_out/bench_ours_strong.log 2>&1; echo "strong exit $?"
tail -1 gpurun_out/bench_ours_strong.log
timeout 300 python bench.py --impl reference --scaling strong --steps 2 > gpurun_out/bench_ref_strong.log 2>&1; echo "ref strong exit $?"
tail -1 gpurun_out/bench_ref_strong.log
timeout 600 python bench.py --steps 5 --warmup 3 --mode persistent --no-small > gpurun_out/bench_ours_persistent.log 2>&1; echo "persistent exit $?"
tail -1 gpurun_out/bench_ours_persistent.log | cut -c1-400
