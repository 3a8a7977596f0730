#!/bin/bash
# Runtime-switch experiments on the GPU box: bench.py under each environment setting.
#   tools/env_sweep.sh out.log "<bench args>" "" "FIN_SUB_MBASES=24" "FIN_SUB_MBASES=24 FIN_SLOTS=4" ...
out=$1; args=$2; shift 2
mkdir -p "$(dirname "$out")"; : > "$out"
for envs in "$@"; do
  env $envs python bench.py $args --no-cpu-baseline --no-build-arm 2>/tmp/env_err.log | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('env [$envs] args [$args]: %.3e k-mers/s device (kernel %.2f ms), e2e %.3e (%.1f ms/step, floor %.1f)' % (d['value'], d['roofline']['kernel_ms'], d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e'].get('d2h_floor_ms_per_step', 0)))
" >> "$out" || tail -3 /tmp/env_err.log >> "$out"
done
cat "$out"
