#!/bin/bash
# DRAM traffic per launch of the dominant kernel for every named shape (bench.py's
# roofline.traffic), and the launch list of the default bench command. One GPU.
# Each capture follows a plain run of the same command that exited 0.
set -u
mkdir -p gpurun_out/ncu
B="python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-configs --no-variants --init tiled"
M="dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum"
run() {   # name, extra bench flags
  local name=$1; shift
  $B "$@" > gpurun_out/ncu/$name.plain.json 2> gpurun_out/ncu/$name.plain.err || { echo "$name: plain run failed"; return; }
  ncu --metrics $M --clock-control none -k regex:"lb_sweep_kernel|phi_from_g_kernel" -c 24 --csv \
      --log-file gpurun_out/ncu/$name.csv $B "$@" > gpurun_out/ncu/$name.ncu.out 2>&1
  echo "$name rc=$?"
}
run q19-256-ondemand --hydro-write 0
run q19-256-store    --hydro-write 1
run q19-64-ondemand  --hydro-write 0 --lsize 64 64 64
run q19-binary-128   --hydro-write 0 --ndist 2 --lsize 128 128 128
run q27-256-ondemand --hydro-write 0 --nvel 27
run q9-4096-ondemand --hydro-write 0 --nvel 9
# launch list of the default command (every kernel, serialised and cold: the SHARE of the
# sweep in a step is what must agree with bench.py)
python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu/default.plain.json 2> gpurun_out/ncu/default.plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/ncu/r2_launches_default.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu/default.ncu.out 2>&1
echo "launch list rc=$?"
# one full capture of the headline sweep with source
ncu --set full --clock-control none --import-source on -k regex:lb_sweep_kernel --launch-skip 4 -c 1 \
    -o gpurun_out/ncu/r2_sweep_q19_256_full $B --hydro-write 0 > gpurun_out/ncu/full.ncu.out 2>&1
echo "full rc=$?"
ncu -i gpurun_out/ncu/r2_sweep_q19_256_full.ncu-rep --page raw --csv > gpurun_out/ncu/r2_sweep_q19_256_full.raw.csv 2>/dev/null
ncu -i gpurun_out/ncu/r2_sweep_q19_256_full.ncu-rep --page details > gpurun_out/ncu/r2_sweep_q19_256_full.details.txt 2>/dev/null
ls -la gpurun_out/ncu | head -40
