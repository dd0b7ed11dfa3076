# round 2, session 12: packed z-face planes on two real GPUs (NVLink): parity, then 1x1x2 with and without
mkdir -p gpurun_out
export B200LB_ASSUME_GPU=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541"
( time timeout 600 python -m pytest tests/test_multi_gpu.py -x -q -m gpu -p no:cacheprovider -k "two_ranks_along_other_axes or two_processes" ) > gpurun_out/r2b_pytest_mgpu2_zpack.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2b_pytest_mgpu2_zpack.log
for z in 1 2 0; do
  timeout 300 $TR bench.py --gpus 2 --grid 1 1 2 --zpack $z --steps 200 --warmup 10 --no-e2e --no-variants --no-configs > gpurun_out/r2b_bench_2gpu_112_zpack$z.json 2> gpurun_out/r2b_bench_2gpu_112_zpack$z.err; echo "112 zpack $z rc=$?"
  python -c "import json;d=json.load(open('gpurun_out/r2b_bench_2gpu_112_zpack$z.json'));print('zpack $z', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['parity'], d['config'].get('z_faces_packed'))"
done
timeout 300 $TR bench.py --gpus 2 --steps 200 --warmup 10 --no-e2e --no-variants --no-configs --no-parity > gpurun_out/r2b_bench_2gpu_211.json 2> gpurun_out/r2b_bench_2gpu_211.err; echo "211 rc=$?"
python -c "import json;d=json.load(open('gpurun_out/r2b_bench_2gpu_211.json'));print('211', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'])"
timeout 200 $TR tests/mgpu_worker.py --nvel 19 --halo reduced --grid 1 1 2 --lsize 16 12 64 --steps 12 > gpurun_out/r2b_mgpu2_112_zpack.log 2>&1; echo "worker rc=$?"; tail -1 gpurun_out/r2b_mgpu2_112_zpack.log
