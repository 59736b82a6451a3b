#!/bin/bash
# End-of-round evidence on one B200 box: the whole GPU suite, the bench lines of all five configurations and the reference
# arm, the ncu launch list of one bench step, and one `ncu --set full` tour of every kernel (reports stay in /tmp on the box:
# they exceed gpurun_out's 64 MiB; their text / csv summaries come back).  A number printed under ncu is never a bench value.
mkdir -p gpurun_out/final
O=gpurun_out/final
timeout 900 python -m pytest tests -m gpu -x -q > $O/gpu_tests.log 2>&1; echo "pytest rc=$?" | tee -a $O/gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench_cfg1.json 2> $O/bench_cfg1.err; echo "bench cfg1 rc=$?"
for c in 0 2 3 4; do
    timeout 900 python bench.py --config $c > $O/bench_cfg$c.json 2> $O/bench_cfg$c.err; echo "bench cfg$c rc=$?"
done
if [ "$1" != "noref" ]; then
    timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > $O/ref_cfg1.json 2> $O/ref_cfg1.err; echo "reference rc=$?"
fi
FPS=${FRAMES_PER_STEP:-16}
B="python bench.py --steps 1 --warmup 1 --frames-per-step $FPS --no-cpu-baseline --profiler-range"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file $O/launches.csv $B > $O/ncu_launches.log 2>&1; echo "ncu-launches rc=$?"
python tools/ncu_summary.py launches $O/launches.csv "configs[1], one step of $FPS frames (bench.py --steps 1 --frames-per-step $FPS --profiler-range)" > $O/ncu_launches.txt 2>&1
gzip -f $O/launches.csv
if [ "$2" = "notour" ]; then tail -2 $O/gpu_tests.log; tail -1 $O/smoke.log; for c in 1 0 2 3 4; do head -c 200 $O/bench_cfg$c.json; echo; done; head -12 $O/ncu_launches.txt; exit 0; fi
timeout 300 python tools/kernel_tour.py > $O/tour_plain.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on --metrics $(python tools/ncu_summary.py extra-metrics) \
    --kernel-id "::regex:.*:2" -f -o /tmp/tour python tools/kernel_tour.py > $O/ncu_tour.log 2>&1; echo "ncu-tour rc=$?"
python tools/ncu_summary.py bounds /tmp/tour.ncu-rep $O/kernel_bounds.json > $O/tour_bounds.txt 2>&1
python tools/ncu_summary.py full /tmp/tour.ncu-rep > $O/tour_full.txt 2>&1
ncu -i /tmp/tour.ncu-rep --page raw --csv 2>/dev/null | gzip > $O/tour_raw.csv.gz
# DRAM traffic of ONE packed matcher call (pack + 2 x (GEMM, eval, merge)): the 7 kernels after the first tour pass
timeout 600 ncu --set full --clock-control none -k regex:feat_tc --launch-skip 21 -c 7 -f -o /tmp/full_feat python tools/kernel_tour.py > $O/ncu_feat.log 2>&1; echo "ncu-feat rc=$?"
python tools/ncu_summary.py traffic feature_nn=/tmp/full_feat.ncu-rep > $O/traffic.txt 2>&1; cp profiles/ncu_traffic.json $O/ncu_traffic.json
tail -2 $O/gpu_tests.log; for c in 1 0 2 3 4; do head -c 300 $O/bench_cfg$c.json; echo; done; ls -la $O | head -40
