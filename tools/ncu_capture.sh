#!/bin/bash
# ncu --set full captures of the search kernels (strand_filter_kernel + search_reads_kernel, first dense launch) on the GPU box:
# configs[1] (5 Mbp, 1 M reads) and the HBM regime (1 Gbp, 4 M reads).  Each capture runs only after the same command exited 0
# without ncu.  Read the .ncu-rep files here with tools/ncu_traffic.py / tools/ncu_lines.py.
#   tools/ncu_capture.sh <tag>
tag=${1:-cap}
mkdir -p gpurun_out
for spec in "5000000 1000000 cfg2" "1000000000 4000000 1gbp"; do
  set -- $spec
  python tools/kernel_probe.py --genome $1 --reads $2 --steps 1 --no-e2e > gpurun_out/${tag}_$3_probe.log 2>&1 || { tail -3 gpurun_out/${tag}_$3_probe.log; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'strand_filter_kernel|search_reads_kernel' -c 2 -f \
      -o gpurun_out/${tag}_search_$3 python tools/kernel_probe.py --genome $1 --reads $2 --steps 1 --no-e2e > gpurun_out/${tag}_$3_ncu.log 2>&1
  tail -1 gpurun_out/${tag}_$3_probe.log | cut -c1-300
done
ls -la gpurun_out | grep ncu-rep
