#!/bin/bash
# Kernel experiments on the GPU box: rebuilds the library with each set of extra nvcc flags and runs tools/kernel_probe.py
# on the L2-resident config (5 Mbp) and the HBM-regime one (200 Mbp).
#   tools/variant_sweep.sh out.log "" "-DFIN_SEARCH_MIN_BLOCKS=10" ...        (GENOMES, READS, PROBE_ARGS override)
out=$1; shift
mkdir -p "$(dirname "$out")"; : > "$out"
for extra in "$@"; do
  touch finito_b200/csrc/search.cu
  make -C finito_b200/csrc all EXTRA="$extra" > /tmp/variant_make.log 2>&1 || { echo "variant [$extra]: build failed" >> "$out"; tail -5 /tmp/variant_make.log >> "$out"; continue; }
  for kern in "search_reads_kernelILi32EiLb0ELb0" "search_reads_kernelILi32EiLb0ELb1"; do
    grep -A2 "$kern" finito_b200/build/search.ptxas.log | grep -E "registers|spill" | tr '\n' ' ' >> "$out"; echo >> "$out"
  done
  for g in ${GENOMES:-5000000 1000000000}; do
    python tools/kernel_probe.py --genome $g --reads ${READS:-2000000} --tag="$extra" $PROBE_ARGS >> "$out" 2>/tmp/variant_err.log || tail -3 /tmp/variant_err.log >> "$out"
  done
done
touch finito_b200/csrc/search.cu; make -C finito_b200/csrc all > /dev/null 2>&1
cat "$out"
