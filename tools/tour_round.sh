#!/bin/bash
# `ncu --set full` tour of every kernel of the library (second invocation of each) and a per-instruction capture of the step's
# top kernels, on the current build; reports stay in /tmp on the box, their summaries come back under gpurun_out/tour/.
mkdir -p gpurun_out/tour
O=gpurun_out/tour
timeout 300 python tools/kernel_tour.py > $O/tour_plain.log 2>&1 || { echo "tour failed"; tail -5 $O/tour_plain.log; exit 1; }
timeout 1200 ncu --set full --clock-control none --import-source on --metrics $(python tools/ncu_summary.py extra-metrics) \
    --kernel-id "::regex:.*:2" -f -o /tmp/tour python tools/kernel_tour.py > $O/ncu_tour.log 2>&1; echo "ncu-tour rc=$?"
python tools/ncu_summary.py bounds /tmp/tour.ncu-rep $O/kernel_bounds.json > $O/tour_bounds.txt 2>&1
python tools/ncu_summary.py full /tmp/tour.ncu-rep > $O/tour_full.txt 2>&1
ncu -i /tmp/tour.ncu-rep --page raw --csv 2>/dev/null | gzip > $O/tour_raw.csv.gz
timeout 600 ncu --set full --clock-control none --import-source on \
  --kernel-id "::regex:knn_kernel|fpfh_kernel|spfh_kernel|feat_tc_eval_kernel|feat_tc_kernel|ransac_count_batch_kernel|ransac_edge_batch_kernel:3" \
  -f -o /tmp/src_top python tools/kernel_tour.py > $O/ncu_src.log 2>&1; echo "ncu-src rc=$?"
ncu -i /tmp/src_top.ncu-rep --page source --csv 2>/dev/null | gzip > $O/src_top_source.csv.gz
ncu -i /tmp/src_top.ncu-rep --page raw --csv 2>/dev/null | gzip > $O/src_top_raw.csv.gz
ls -la $O; head -25 $O/tour_bounds.txt
