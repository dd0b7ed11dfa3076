mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -q -p no:cacheprovider --timeout 600 -k "lattices or two_ranks or two_processes" > gpurun_out/r2_pytest_mgpu2_aa.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2_pytest_mgpu2_aa.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541"
timeout 300 $TR bench.py --gpus 2 --steps 200 --warmup 5 --no-e2e --init tiled --lattices 1 > gpurun_out/r2_bench_2gpu_aa_256.json 2> gpurun_out/r2_bench_2gpu_aa.err; echo "256 rc=$?"
timeout 1200 $TR bench.py --gpus 2 --steps 20 --warmup 3 --no-e2e --no-parity --lattices 1 --global-size 1024 1024 1024 > gpurun_out/r2_bench_2gpu_strong1024_aa.json 2>> gpurun_out/r2_bench_2gpu_aa.err; echo "1024 rc=$?"
grep -v "OMP_NUM_THREADS\|\*\*\*\*" gpurun_out/r2_bench_2gpu_aa.err | tail -12
