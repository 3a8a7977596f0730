#!/bin/bash
# The record run of a round on the GPU box: parity tests, smoke, the default bench (configs[1] + the HBM-regime entry), the
# reference arm, and the k=63 / 100 Mbp config.  Outputs under gpurun_out/<tag>_*.
tag=${1:-rec}
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_pytest.log)
timeout 300 python __graft_entry__.py smoke > gpurun_out/${tag}_smoke.log 2>&1
timeout 900 python bench.py > gpurun_out/${tag}_bench.log 2> gpurun_out/${tag}_bench.err
timeout 400 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/${tag}_bench_ref.log 2>&1
timeout 900 python bench.py --genome 100000000 --k 63 --reads 5000000 --no-cpu-baseline --no-hbm-arm > gpurun_out/${tag}_bench_c4_k63.log 2>&1
tail -3 gpurun_out/${tag}_pytest.log; tail -1 gpurun_out/${tag}_smoke.log; tail -c 300 gpurun_out/${tag}_bench.err
for f in bench bench_ref bench_c4_k63; do tail -1 gpurun_out/${tag}_$f.log | cut -c1-260; done
