mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29531"
timeout 400 $TR bench.py --gpus 4 --steps 200 --warmup 5 --no-e2e --init tiled > gpurun_out/r2_bench_4gpu_221.json 2> gpurun_out/r2_bench_4gpu.err; echo "221 rc=$?"
timeout 600 $TR bench.py --gpus 4 --steps 50 --warmup 3 --no-e2e --no-parity --lsize 512 512 512 > gpurun_out/r2_bench_4gpu_221_512.json 2>> gpurun_out/r2_bench_4gpu.err; echo "512 weak rc=$?"
timeout 900 $TR bench.py --gpus 4 --steps 30 --warmup 3 --no-e2e --no-parity --global-size 1024 1024 1024 > gpurun_out/r2_bench_4gpu_strong1024.json 2>> gpurun_out/r2_bench_4gpu.err; echo "1024 strong rc=$?"
grep -v "OMP_NUM_THREADS\|\*\*\*\*" gpurun_out/r2_bench_4gpu.err | tail -8
