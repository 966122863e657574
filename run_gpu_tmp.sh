python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -5 gpurun_out/pytest_gpu.log
python bench.py --no-cpu-baseline > gpurun_out/bench_default.log 2>&1; tail -c 1800 gpurun_out/bench_default.log
