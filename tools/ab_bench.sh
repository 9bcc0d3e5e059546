#!/bin/bash
# A/B of bench.py under environment toggles: tools/ab_bench.sh "VAR=a VAR=b ..." [steps...]
# prints ms_per_step, the run kernel's launch_ms and the host-buffer e2e ms for every (setting, steps) pair
settings="$1"; shift
steps="${@:-20 50}"
for s in $settings; do
  for st in $steps; do
    env $s python bench.py --no-cpu --no-extra --steps $st 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$s', 'steps', $st, 'ms_per_step %.4f' % d['ms_per_step'], 'launch_ms %.4f' % d['roofline']['launch_ms'], 'e2e_ms %.4f' % d['e2e']['ms_per_step'], 'rebuilds/slice', d['config'].get('rebuilds_per_slice'))
"
  done
done
