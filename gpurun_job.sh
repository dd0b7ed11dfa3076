#!/bin/bash
# the whole GPU suite (with the viscosity-model tests) and a short default bench line
mkdir -p gpurun_out
timeout 150 python -m pytest tests -q -m gpu -p no:cacheprovider --tb=short > gpurun_out/t_full.log 2>&1; echo "rc=$?" >> gpurun_out/t_full.log
tail -25 gpurun_out/t_full.log
timeout 60 python bench.py --steps 100 --no-cpu --no-e2e > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; tail -2 gpurun_out/bench_quick.err; cut -c1-330 gpurun_out/bench_quick.json
