mkdir -p gpurun_out/sanitizer
S=gpurun_out/sanitizer
python tools/p2p_two_process.py > $S/plain_112.log 2>&1; echo "plain 112 rc=$?"
python tools/p2p_two_process.py --grid 2 1 1 > $S/plain_211.log 2>&1; echo "plain 211 rc=$?"
timeout 600 compute-sanitizer --tool memcheck --target-processes all --print-limit 20 python tools/p2p_two_process.py > $S/memcheck_2gpu_112.log 2>&1; echo "memcheck 2gpu rc=$?"
timeout 600 compute-sanitizer --tool racecheck --target-processes all --print-limit 20 python tools/p2p_two_process.py > $S/racecheck_2gpu_112.log 2>&1; echo "racecheck 2gpu rc=$?"
timeout 600 compute-sanitizer --tool memcheck --print-limit 20 python tools/p2p_two_process.py --one-gpu > $S/memcheck_1gpu_loopback.log 2>&1; echo "memcheck 1gpu rc=$?"
timeout 600 compute-sanitizer --tool racecheck --print-limit 20 python tools/p2p_two_process.py --one-gpu > $S/racecheck_1gpu_loopback.log 2>&1; echo "racecheck 1gpu rc=$?"
timeout 600 compute-sanitizer --tool synccheck --print-limit 20 python tools/p2p_two_process.py --one-gpu > $S/synccheck_1gpu_loopback.log 2>&1; echo "synccheck 1gpu rc=$?"
tail -4 $S/*.log
