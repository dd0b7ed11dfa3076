mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521"
timeout 400 $TR bench.py --gpus 8 --steps 200 --warmup 5 --no-e2e --init tiled > gpurun_out/r2_bench_8gpu_222.json 2> gpurun_out/r2_bench_8gpu.err; echo "222 rc=$?"
timeout 300 $TR bench.py --gpus 8 --steps 200 --warmup 5 --no-e2e --no-parity --init tiled --grid 4 2 1 > gpurun_out/r2_bench_8gpu_421.json 2>> gpurun_out/r2_bench_8gpu.err; echo "421 rc=$?"
timeout 500 $TR bench.py --gpus 8 --steps 50 --warmup 3 --no-e2e --no-parity --local 512 512 512 > gpurun_out/r2_bench_8gpu_222_512.json 2>> gpurun_out/r2_bench_8gpu.err; echo "512 rc=$?"
timeout 200 $TR tests/mgpu_worker.py --nvel 19 --halo reduced > gpurun_out/r2_mgpu8_reduced.log 2>&1; echo "w1 rc=$?"
timeout 200 $TR tests/mgpu_worker.py --nvel 19 --ndist 2 --halo full --lsize 8 8 6 --gradient 27 > gpurun_out/r2_mgpu8_binary27.log 2>&1; echo "w2 rc=$?"
tail -2 gpurun_out/r2_mgpu8_reduced.log gpurun_out/r2_mgpu8_binary27.log
grep -v "OMP_NUM_THREADS\|\*\*\*\*" gpurun_out/r2_bench_8gpu.err | tail -5
