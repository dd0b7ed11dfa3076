# round 2, session 12: everything the driver runs at round end, at HEAD, on one B200
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider --timeout 900 ) > gpurun_out/r2b_pytest_gpu_full.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2b_pytest_gpu_full.log
( time python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r2b_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/r2b_smoke.log
( time python bench.py --gpus 1 --steps 20 --warmup 3 ) > gpurun_out/r2b_bench_default_20.json 2> gpurun_out/r2b_bench_default_20.err; echo "bench20 rc=$?"; tail -3 gpurun_out/r2b_bench_default_20.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2b_bench_default_20.json') if l.startswith('{')][-1])
print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['gpu_launches'], d['clocks'])
for c in d.get('configs') or []: print(c.get('workload', c.get('config')), c.get('value'), (c.get('roofline') or {}).get('step_frac'))
PY
( time python bench.py --loopback 15 --halo reduced --steps 200 --warmup 5 --no-cpu --no-e2e --no-configs --no-variants ) > gpurun_out/r2b_bench_loopback_xyz.json 2> gpurun_out/r2b_bench_loopback_xyz.err; echo "loopback rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2b_bench_loopback_xyz.json') if l.startswith('{')][-1])
print('loopback xyz', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['config'].get('z_faces_packed'), d['launches_per_step'])
PY
