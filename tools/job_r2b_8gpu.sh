# round 2, session 12: 2x2x2 at 256^3 per GPU with the packed z-face planes (8 GPUs)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551"
timeout 300 $TR bench.py --gpus 8 --steps 200 --warmup 5 --no-e2e --no-variants --init tiled > gpurun_out/r2b_bench_8gpu_222.json 2> gpurun_out/r2b_bench_8gpu_222.err; echo "222 rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2b_bench_8gpu_222.json') if l.startswith('{')][-1])
print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['parity'], d['config'].get('z_faces_packed'), d['config'].get('decomposition'))
PY
grep -v "OMP_NUM_THREADS\|\*\*\*\*" gpurun_out/r2b_bench_8gpu_222.err | tail -3
