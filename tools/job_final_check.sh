mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider --timeout 900 ) > gpurun_out/r2_pytest_gpu_full.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r2_pytest_gpu_full.log
( time python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r2_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/r2_smoke.log
( time python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 ) > gpurun_out/r2_bench_ref_default.json 2> gpurun_out/r2_bench_ref_default.err; echo "ref rc=$?"; tail -3 gpurun_out/r2_bench_ref_default.err
( time python bench.py --gpus 1 --steps 20 --warmup 3 ) > gpurun_out/r2_bench_default_20.json 2> gpurun_out/r2_bench_default_20.err; echo "bench20 rc=$?"; tail -3 gpurun_out/r2_bench_default_20.err
( time python bench.py ) > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; echo "bench rc=$?"; tail -3 gpurun_out/r2_bench_default.err
