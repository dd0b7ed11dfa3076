# round 2, session 12: packed z-face planes, grad27 prefetch, binary block shapes (one GPU)
mkdir -p gpurun_out
export B200LB_ASSUME_GPU=1
( time timeout 600 python -m pytest tests/test_zzzz_gpu_zpack.py tests/test_zzzzz_no_alloc.py -x -q -m gpu -p no:cacheprovider ) > gpurun_out/r2b_pytest_zpack.log 2>&1; echo "pytest zpack rc=$?"; tail -5 gpurun_out/r2b_pytest_zpack.log
( time timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -p no:cacheprovider -k "loopback or peer or grad" ) > gpurun_out/r2b_pytest_loopback.log 2>&1; echo "pytest loopback rc=$?"; tail -3 gpurun_out/r2b_pytest_loopback.log
timeout 600 python tools/exp_push_axes.py --size 256 --steps 100 > gpurun_out/r2b_push_axes_zpack.log 2>&1; echo "push axes rc=$?"; cat gpurun_out/r2b_push_axes_zpack.log
timeout 300 python tools/time_aux_kernels.py > gpurun_out/r2b_aux_kernels_256.log 2>&1; echo "aux rc=$?"; cat gpurun_out/r2b_aux_kernels_256.log
for shape in "32 128" "64 128" "128 128" "32 256" "64 256" "128 256" "64 64" "32 64"; do
  set -- $shape
  tag="b$1_t$2"
  timeout 300 python bench.py --ndist 2 --local 128 128 128 --block $1 --threads $2 --steps 200 --warmup 20 --no-cpu --no-e2e --no-configs --no-variants > gpurun_out/r2b_binary128_$tag.json 2> gpurun_out/r2b_binary128_$tag.err; echo "binary $tag rc=$?"
  python -c "import json;d=json.load(open('gpurun_out/r2b_binary128_$tag.json'));print('$tag', d['value'], d['ms_per_step'], d['roofline'])"
done
timeout 300 python bench.py --ndist 2 --local 256 256 256 --steps 50 --warmup 5 --no-cpu --no-e2e --no-configs --no-variants > gpurun_out/r2b_binary256.json 2> gpurun_out/r2b_binary256.err; echo "binary 256 rc=$?"
python -c "import json;d=json.load(open('gpurun_out/r2b_binary256.json'));print('binary256', d['value'], d['ms_per_step'], d['roofline'])"
timeout 300 python bench.py --local 128 128 128 --steps 200 --warmup 20 --no-cpu --no-e2e --no-configs --no-variants > gpurun_out/r2b_q19_128.json 2> gpurun_out/r2b_q19_128.err; echo "q19 128 rc=$?"
python -c "import json;d=json.load(open('gpurun_out/r2b_q19_128.json'));print(d['value'], d['ms_per_step'], d['roofline'])"
