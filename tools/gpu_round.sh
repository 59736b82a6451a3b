#!/bin/bash
# One GPU round trip: parity tests, bench, ncu launch list of the timed region, ncu --set full of the top kernels.
# Every stage has its own timeout; a number printed under ncu is never a bench value.
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --frames-per-step 1 --workers 1 --no-cpu-baseline --profiler-range"
timeout 700 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gpu_tests.log
timeout 420 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" >> gpurun_out/bench.err
timeout 420 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launches.log 2>&1; echo "ncu-launches rc=$?" >> gpurun_out/ncu_launches.log
# feat_tc: the 10 kernels of ONE ibl_feature_nn call (centre, range x2, pack x2, rows32 x2, one-pass GEMM, eval, merge), a 3-object call
for spec in "feat_tc:80:10" "ransac_count_kernel:6:2" "ransac_score_kernel:6:2" "colored_step_kernel:10:2" "knn_kernel:10:2" "fpfh_kernel:5:1"; do
    IFS=: read k skip cnt <<< "$spec"
    timeout 300 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:$k \
        --launch-skip $skip -c $cnt -f -o gpurun_out/full_$k $B > gpurun_out/ncu_full_$k.log 2>&1
    echo "ncu-full $k rc=$?" >> gpurun_out/ncu_launches.log
done
# memory-build / input kernels (outside the bench's timed region): CUDA-event probes, then one ncu --set full capture each
timeout 120 python tools/dbscan_probe.py 200000 > gpurun_out/dbscan_probe.jsonl 2> gpurun_out/dbscan_probe.err
timeout 60 python tools/raycast_probe.py > gpurun_out/raycast_probe.jsonl 2> gpurun_out/raycast_probe.err
timeout 300 ncu --set full --clock-control none --import-source on -k regex:dbscan_walk_kernel --launch-skip 2 -c 2 -f \
    -o gpurun_out/full_dbscan_walk python tools/dbscan_probe.py 200000 > gpurun_out/ncu_full_dbscan.log 2>&1
timeout 120 ncu --set full --clock-control none --import-source on -k regex:raycast_kernel --launch-skip 1 -c 1 -f \
    -o gpurun_out/full_raycast python tools/raycast_probe.py > gpurun_out/ncu_full_raycast.log 2>&1
tail -3 gpurun_out/gpu_tests.log; head -c 400 gpurun_out/bench.json; echo; tail -3 gpurun_out/bench.err; tail -6 gpurun_out/ncu_launches.log
