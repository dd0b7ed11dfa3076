#!/bin/bash
# one short call: the finite-difference order-parameter kernels (tests, host mirror,
# regression logs, one bench line)
mkdir -p gpurun_out
export B200LB_ASSUME_GPU=1
nvidia-smi -L > gpurun_out/gpu.txt 2>&1
timeout 170 python -m pytest tests/test_zz_gpu_cahn_hilliard.py -q -m gpu --maxfail=8 --tb=short -p no:cacheprovider > gpurun_out/t_ch.log 2>&1; echo "rc=$?" >> gpurun_out/t_ch.log
tail -30 gpurun_out/t_ch.log
timeout 60 tests/host/test_host_shim.exe fd > gpurun_out/t_shim_fd.log 2>&1; echo "rc=$?" >> gpurun_out/t_shim_fd.log
tail -6 gpurun_out/t_shim_fd.log
timeout 150 python -m pytest tests/test_regression_logs.py -q -m gpu -k "fd1 or dr1 or pat or muex" --tb=short -p no:cacheprovider > gpurun_out/t_reg_fd.log 2>&1; echo "rc=$?" >> gpurun_out/t_reg_fd.log
tail -12 gpurun_out/t_reg_fd.log
timeout 60 python tools/run_fd_steps.py 256 1 27 20 > gpurun_out/fd_steps.log 2>&1; cat gpurun_out/fd_steps.log
timeout 90 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 40 --csv --log-file gpurun_out/fd_launches.csv python tools/run_fd_steps.py 256 1 27 2 > gpurun_out/fd_ncu.log 2>&1; tail -2 gpurun_out/fd_ncu.log
timeout 150 python bench.py --fd 1 --gradient 27 --steps 50 --warmup 3 --no-cpu --no-e2e > gpurun_out/bench_fd1.json 2> gpurun_out/bench_fd1.err; tail -3 gpurun_out/bench_fd1.err; cat gpurun_out/bench_fd1.json
timeout 100 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "binary or gradient or time_loop" -p no:cacheprovider > gpurun_out/t_sanity.log 2>&1; echo "rc=$?" >> gpurun_out/t_sanity.log
tail -4 gpurun_out/t_sanity.log
