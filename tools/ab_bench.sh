#!/bin/bash
# A/B of bench.py under environment switches:  tools/ab_bench.sh <steps> "<label>=<ENV=1 ...>" ...
STEPS=$1; shift
for spec in "$@"; do
  label=${spec%%=*}; envs=${spec#*=}
  [ "$envs" = "$spec" ] && envs=""
  env $envs timeout 300 python bench.py --steps $STEPS --no-cpu-baseline 2>gpurun_out/ab_$label.err > gpurun_out/ab_$label.json
  python - "$label" <<'PY'
import json,sys
lab=sys.argv[1]
try:
    d=json.load(open('gpurun_out/ab_%s.json'%lab))
    print("%-14s value %8.1f  e2e %8.1f  equal %s  wall/step %.1f us  stage %s" % (lab, d['value'], d['e2e']['value'], d.get('pipelined_poses_equal_synchronous'), 1e3*d.get('wall_ms_per_step_incl_flush',0), {k: round(v*1e3,1) for k,v in d['stage_ms'].items()}))
except Exception as e:
    print(lab, "FAILED", e); print(open('gpurun_out/ab_%s.err'%lab).read()[-1500:])
PY
done
