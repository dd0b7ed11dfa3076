# round 2, session 12, last call: the driver's sequence at HEAD, and one full ncu capture of the peer-memory sweep
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider --timeout 900 ) > gpurun_out/r2b_pytest_gpu_full4.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2b_pytest_gpu_full4.log
( time python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r2b_smoke4.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2b_smoke4.log
timeout 200 ncu --set full --clock-control none --import-source on -k regex:lb_sweep_kernel -s 6 -c 1 -o gpurun_out/r2b_push_sweep_full -f python bench.py --loopback 15 --halo reduced --steps 3 --warmup 3 --no-cpu --no-e2e --no-configs --no-variants > gpurun_out/r2b_ncu_full.log 2>&1; echo "ncu rc=$?"; ls -la gpurun_out/r2b_push_sweep_full.ncu-rep
