#!/bin/bash
mkdir -p gpurun_out
export B200LB_ASSUME_GPU=1
timeout 9 python -m pytest tests/test_zzz_gpu_g_moment.py -x -q -m gpu > gpurun_out/gmoment_pytest2.log 2>&1; echo "rc=$?" >> gpurun_out/gmoment_pytest2.log
tail -5 gpurun_out/gmoment_pytest2.log
