mkdir -p gpurun_out
python tools/exp_push_axes.py > gpurun_out/r2_push_axes_pdl.log 2>&1; cat gpurun_out/r2_push_axes_pdl.log
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -p no:cacheprovider --timeout 300 -k "loopback or peer" > gpurun_out/r2_pytest_pdl_1gpu.log 2>&1; echo "1gpu pytest rc=$?"; tail -3 gpurun_out/r2_pytest_pdl_1gpu.log
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -q -p no:cacheprovider --timeout 600 -k "two_ranks or two_processes or extra1-2 or extra7-2 or extra8-2 or extra10-2 or extra25-2 or extra26-2 or extra27-2" > gpurun_out/r2_pytest_mgpu2_pdl.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2_pytest_mgpu2_pdl.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551"
for g in "2 1 1" "1 1 2"; do timeout 300 $TR bench.py --gpus 2 --steps 200 --warmup 5 --no-e2e --init tiled --grid $g > "gpurun_out/r2_bench_2gpu_pdl_$(echo $g | tr -d " ").json" 2> gpurun_out/r2_bench_2gpu_pdl.err; echo "grid $g rc=$?"; done
grep -v "OMP_NUM_THREADS\|\*\*\*\*" gpurun_out/r2_bench_2gpu_pdl.err | tail -5
