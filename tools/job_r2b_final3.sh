# round 2, session 12: full suite and both bench forms after stats_final became two kernels; launch list for its timing
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider --timeout 900 ) > gpurun_out/r2b_pytest_gpu_full3.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2b_pytest_gpu_full3.log
( time python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r2b_smoke3.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2b_smoke3.log
( time python bench.py ) > gpurun_out/r2b_bench_default3.json 2> gpurun_out/r2b_bench_default3.err; echo "bench rc=$?"; tail -3 gpurun_out/r2b_bench_default3.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2b_bench_default3.json') if l.startswith('{')][-1])
print(d['steps'], d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['e2e']['loop_seconds'], d['e2e']['copy_seconds'], d['gpu_launches'], d['clocks'])
PY
( time python bench.py --steps 20 --warmup 3 --no-configs --no-variants --no-cpu ) > gpurun_out/r2b_bench_20_3.json 2> gpurun_out/r2b_bench_20_3.err; echo "bench20 rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2b_bench_20_3.json') if l.startswith('{')][-1])
print(d['steps'], d['value'], d['e2e']['value'], d['e2e']['loop_seconds'], d['e2e']['copy_seconds'])
PY
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:stats -c 12 --csv --log-file gpurun_out/r2b_ncu_stats.csv python bench.py --steps 3 --warmup 3 --no-cpu --no-configs --no-variants > gpurun_out/r2b_ncu_stats.log 2>&1; echo "ncu rc=$?"; grep -c stats gpurun_out/r2b_ncu_stats.csv
