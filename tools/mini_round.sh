mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --frames-per-step 1 --workers 1 --no-cpu-baseline --profiler-range"
timeout 420 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launches.log 2>&1; echo "ncu-launches rc=$?"
timeout 300 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:feat_tc --launch-skip 80 -c 10 -f -o gpurun_out/full_feat_tc $B > gpurun_out/ncu_full_feat_tc.log 2>&1; echo "ncu-full rc=$?"
timeout 200 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; head -c 150 gpurun_out/bench.json
