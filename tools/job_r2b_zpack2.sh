# round 2, session 12: the four arrangements of the packed z-face planes; reads and writes timed apart
mkdir -p gpurun_out
export B200LB_ASSUME_GPU=1
( time timeout 600 python -m pytest tests/test_zzzz_gpu_zpack.py -x -q -m gpu -p no:cacheprovider ) > gpurun_out/r2b_pytest_zpack2.log 2>&1; echo "pytest zpack rc=$?"; tail -5 gpurun_out/r2b_pytest_zpack2.log
timeout 600 python tools/exp_push_axes.py --size 256 --steps 100 > gpurun_out/r2b_push_axes_zpack2.log 2>&1; echo "push axes rc=$?"; cut -c1-200 gpurun_out/r2b_push_axes_zpack2.log
