#!/bin/bash
# round-1 final: the whole GPU suite, the bench lines, smoke
mkdir -p gpurun_out
timeout 160 python -m pytest tests -q -m gpu -p no:cacheprovider > gpurun_out/t_full.log 2>&1; echo "rc=$?" >> gpurun_out/t_full.log
tail -6 gpurun_out/t_full.log
timeout 50 python bench.py --fd 1 --gradient 27 --steps 100 --no-cpu > gpurun_out/bench_fd1.json 2> gpurun_out/bench_fd1.err; tail -2 gpurun_out/bench_fd1.err; cat gpurun_out/bench_fd1.json
timeout 30 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 90 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; tail -2 gpurun_out/bench_default.err; cat gpurun_out/bench_default.json
