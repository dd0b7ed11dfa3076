mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521"
timeout 600 $TR bench.py --gpus 8 --steps 50 --warmup 3 --no-e2e --no-parity --lsize 512 512 512 > gpurun_out/r2_bench_8gpu_222_512.json 2> gpurun_out/r2_bench_8gpu_512.err; echo "512 rc=$?"
grep -v "OMP_NUM_THREADS\|\*\*\*\*" gpurun_out/r2_bench_8gpu_512.err | tail -8
