timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/bench_lr2.json 2> gpurun_out/bench_lr2.err; tail -c 300 gpurun_out/bench_lr2.err
