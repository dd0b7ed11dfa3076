# round 2, session 12: full suite after the stats_final_kernel change, both bench forms, ncu launch list of the push sweep
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider --timeout 900 ) > gpurun_out/r2b_pytest_gpu_full2.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2b_pytest_gpu_full2.log
( time python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r2b_smoke2.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2b_smoke2.log
( time python bench.py ) > gpurun_out/r2b_bench_default.json 2> gpurun_out/r2b_bench_default.err; echo "bench rc=$?"; tail -3 gpurun_out/r2b_bench_default.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2b_bench_default.json') if l.startswith('{')][-1])
print(d['steps'], d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e'], d['gpu_launches'], d['clocks'])
PY
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2b_ncu_push_loopback.csv python bench.py --loopback 15 --halo reduced --steps 3 --warmup 3 --no-cpu --no-e2e --no-configs --no-variants > gpurun_out/r2b_ncu_push_loopback.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/r2b_ncu_push_loopback.log | cut -c1-300
