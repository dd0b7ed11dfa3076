#!/bin/bash
mkdir -p gpurun_out
export B200LB_ASSUME_GPU=1
timeout 100 python -m pytest tests/test_regression_logs.py -q -m gpu -k "fd1 or dr1 or pat or muex" --tb=short -p no:cacheprovider > gpurun_out/t_reg_fd.log 2>&1; echo "rc=$?" >> gpurun_out/t_reg_fd.log
tail -5 gpurun_out/t_reg_fd.log
for pl in 1 2 4 8 16 32 256; do timeout 40 python tools/run_fd_steps.py 256 1 27 20 $pl; done > gpurun_out/fd_planes.log 2>&1
timeout 40 python tools/run_fd_steps.py 256 3 7 20 8 >> gpurun_out/fd_planes.log 2>&1
cat gpurun_out/fd_planes.log
timeout 90 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"ch_kernel|grad27|lb_sweep" -c 15 --csv --log-file gpurun_out/fd_launches.csv python tools/run_fd_steps.py 256 1 27 2 > gpurun_out/fd_ncu.log 2>&1; tail -2 gpurun_out/fd_ncu.log
timeout 120 ncu --set full --clock-control none --import-source on -k regex:ch_kernel -s 3 -c 1 -o gpurun_out/r1_ch_kernel python tools/run_fd_steps.py 256 1 27 2 > gpurun_out/fd_ncu_full.log 2>&1; tail -3 gpurun_out/fd_ncu_full.log
ls -la gpurun_out
