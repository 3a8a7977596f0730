#!/bin/bash
# Kernel experiments on the GPU box: rebuilds the search kernels with each set of extra nvcc flags and runs the
# device arm of bench.py on the L2-resident config (5 Mbp) and the HBM-regime one (200 Mbp).
#   tools/variant_sweep.sh out.log "" "-DFIN_SEARCH_MIN_BLOCKS=10" ...
out=$1; shift
mkdir -p "$(dirname "$out")"; : > "$out"
for extra in "$@"; do
  touch finito_b200/csrc/search.cu
  make -C finito_b200/csrc all EXTRA="$extra" > /tmp/variant_make.log 2>&1 || { echo "variant [$extra]: build failed" >> "$out"; tail -5 /tmp/variant_make.log >> "$out"; continue; }
  grep -A2 "search_reads_kernelILi32EiLb0" finito_b200/build/search.ptxas.log | grep -E "registers|spill" | tr '\n' ' ' >> "$out"; echo >> "$out"
  for g in ${GENOMES:-5000000 200000000}; do
    python bench.py --genome $g --reads ${READS:-2000000} --steps 3 --warmup 2 --no-cpu-baseline --no-build-arm 2>/tmp/variant_err.log | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('variant [$extra] genome $g: %.3e k-mers/s device, kernel %.2f ms, e2e %.3e' % (d['value'], d['roofline']['kernel_ms'], d['e2e']['value']))
" >> "$out" || tail -3 /tmp/variant_err.log >> "$out"
  done
done
touch finito_b200/csrc/search.cu; make -C finito_b200/csrc all > /dev/null 2>&1
cat "$out"
