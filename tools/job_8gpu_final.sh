mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521"
timeout 400 $TR bench.py --gpus 8 --steps 200 --warmup 5 --no-e2e --init tiled > gpurun_out/r2_final_8gpu_222.json 2> gpurun_out/r2_final_8gpu.err; echo "222 rc=$?"
timeout 300 $TR bench.py --gpus 8 --steps 200 --warmup 5 --no-e2e --no-parity --no-variants --init tiled --grid 4 2 1 > gpurun_out/r2_final_8gpu_421.json 2>> gpurun_out/r2_final_8gpu.err; echo "421 rc=$?"
timeout 200 $TR tests/mgpu_worker.py --nvel 19 --halo reduced --lattices 1 --steps 6 > gpurun_out/r2_final_mgpu8_single_lattice.log 2>&1; echo "aa rc=$?"; tail -1 gpurun_out/r2_final_mgpu8_single_lattice.log
timeout 200 $TR tests/mgpu_worker.py --nvel 27 --halo reduced --lsize 8 6 10 > gpurun_out/r2_final_mgpu8_d3q27.log 2>&1; echo "q27 rc=$?"; tail -1 gpurun_out/r2_final_mgpu8_d3q27.log
grep -v "OMP_NUM_THREADS\|\*\*\*\*" gpurun_out/r2_final_8gpu.err | tail -5
