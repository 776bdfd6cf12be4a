#!/bin/bash
# Sweep one environment knob over the sponza stage profile:  tools/sweep_env.sh [LJ_LIB=path] VAR v1 v2 ...
# One line per value in gpurun_out/sweep_<VAR>.txt (and on stdout).
mkdir -p gpurun_out
pre=()
while [[ "$1" == *=* ]]; do pre+=("$1"); shift; done
var=$1; shift
out=gpurun_out/sweep_$var.txt; : > $out
for v in "$@"; do
  env "${pre[@]}" $var=$v python tools/profile_stages.py sponza 64 2>&1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); s=d['stage_ms']
print('$var=$v  Mrays/s %.1f  render %.1f ms  extend %.1f shade %.1f shadow %.1f  nodes/ray %.2f prims/ray %.2f  shadow %.2f/%.2f  bvh nodes %d build %.1f ms' % (d['mrays_per_s'], d['render_ms_plain'], s['extend'], s['shade'], s['shadow+finish'], d['closest']['nodes_per_ray'], d['closest']['prims_per_ray'], d['shadow']['nodes_per_ray'], d['shadow']['prims_per_ray'], d['bvh']['nodes'], d['bvh']['build_ms']))" | tee -a $out
done
