#!/bin/bash
# The record run of a round on the GPU box: parity tests, smoke, the default bench (configs[1]), the reference arm, the
# k=63 / 100 Mbp config, an ncu launch list and one --set full capture of the search kernel.  Outputs under gpurun_out/<tag>_*.
tag=${1:-rec}
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log)
timeout 300 python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1
timeout 900 python bench.py > gpurun_out/${tag}_bench.log 2>&1
timeout 400 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/${tag}_bench_ref.log 2>&1
timeout 900 python bench.py --genome 100000000 --k 63 --reads 5000000 --no-cpu-baseline > gpurun_out/${tag}_bench_c4_k63.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --reads 1000000 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${tag}_ncu_launch.log 2>&1
timeout 500 ncu --set full --clock-control none --import-source on -k regex:search_reads_kernel -s 1 -c 1 -f -o gpurun_out/${tag}_search python bench.py --reads 1000000 --steps 1 --warmup 1 --no-cpu-baseline --no-build-arm > gpurun_out/${tag}_ncu_full.log 2>&1
tail -3 gpurun_out/${tag}_pytest.log; tail -1 gpurun_out/${tag}_smoke.log
for f in bench bench_c4_k63; do tail -1 gpurun_out/${tag}_$f.log | cut -c1-300; done
