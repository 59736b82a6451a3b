#!/bin/bash
# One `ncu --set full --import-source on` capture of the third invocation of each of the step's top kernels (tools/kernel_tour.py,
# configs[1]-sized data), written to /tmp on the box (the report exceeds gpurun_out's 64 MiB), and its per-instruction source
# page + raw metrics brought back as gzipped csv.  Run only after `python tools/kernel_tour.py` has exited 0 without ncu.
mkdir -p gpurun_out
set -x
timeout 300 python tools/kernel_tour.py > gpurun_out/tour_plain.log 2>&1 || { echo "tour failed"; tail -5 gpurun_out/tour_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on \
  --kernel-id "::regex:knn_kernel|fpfh_kernel|spfh_kernel|feat_tc_eval_kernel|feat_tc_kernel|ransac_count_batch_kernel|ransac_check_batch_kernel:3" \
  -f -o /tmp/src_top python tools/kernel_tour.py > gpurun_out/ncu_src.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/src_top.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/src_top_source.csv.gz
ncu -i /tmp/src_top.ncu-rep --page raw --csv 2>/dev/null | gzip > gpurun_out/src_top_raw.csv.gz
ls -la gpurun_out/
