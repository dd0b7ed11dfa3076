#!/bin/bash
mkdir -p gpurun_out
export B200LB_ASSUME_GPU=1
timeout 170 python -m pytest tests/test_zz_gpu_cahn_hilliard.py -q -m gpu --maxfail=8 --tb=short -p no:cacheprovider > gpurun_out/t_ch.log 2>&1; echo "rc=$?" >> gpurun_out/t_ch.log
tail -25 gpurun_out/t_ch.log
timeout 60 tests/host/test_host_shim.exe fd > gpurun_out/t_shim_fd.log 2>&1; echo "rc=$?" >> gpurun_out/t_shim_fd.log
tail -5 gpurun_out/t_shim_fd.log
timeout 100 python -m pytest tests/test_regression_logs.py -q -m gpu -k "fd1 or dr1 or pat or muex" --tb=short -p no:cacheprovider > gpurun_out/t_reg_fd.log 2>&1; echo "rc=$?" >> gpurun_out/t_reg_fd.log
tail -5 gpurun_out/t_reg_fd.log
( timeout 40 python tools/run_fd_steps.py 256 1 27 20 8 0
  timeout 40 python tools/run_fd_steps.py 256 1 27 20 8 1
  timeout 40 python tools/run_fd_steps.py 256 1 27 20 2 0
  timeout 40 python tools/run_fd_steps.py 256 1 27 20 32 0
  timeout 40 python tools/run_fd_steps.py 256 3 7 20 8 0
  timeout 40 python tools/run_fd_steps.py 256 2 7 20 8 0 ) > gpurun_out/fd_variants.log 2>&1
cat gpurun_out/fd_variants.log
timeout 90 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"ch_kernel|grad27|lb_sweep" -c 15 --csv --log-file gpurun_out/fd_launches_fast.csv python tools/run_fd_steps.py 256 1 27 2 > gpurun_out/fd_ncu.log 2>&1; tail -2 gpurun_out/fd_ncu.log
timeout 120 ncu --set full --clock-control none --import-source on -k regex:ch_kernel -s 3 -c 1 -o gpurun_out/r1_ch_kernel_fast python tools/run_fd_steps.py 256 1 27 2 > gpurun_out/fd_ncu_full.log 2>&1; tail -3 gpurun_out/fd_ncu_full.log
