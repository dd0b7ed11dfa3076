#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -q -m gpu -k "random_call" > gpurun_out/t_new.log 2>&1; echo "rc=$?" >> gpurun_out/t_new.log
tail -40 gpurun_out/t_new.log
