#!/bin/bash
# Sweep the trace-kernel knobs (phase-vote threshold, refill threshold, grid size) on sponza 64 spp.
mkdir -p gpurun_out
out=gpurun_out/tune.txt; : > $out
run() { echo "== $*" >> $out; env "$@" python tools/profile_stages.py sponza 64 2>&1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); s=d['stage_ms']
print('Mrays/s %.1f  render_ms %.1f  extend %.1f shade %.1f shadow %.1f gen %.1f  nodes/ray %.2f prims/ray %.2f  sh %.2f/%.2f' % (d['mrays_per_s'], d['render_ms_plain'], s['extend'], s['shade'], s['shadow+finish'], s['generate'], d['closest']['nodes_per_ray'], d['closest']['prims_per_ray'], d['shadow']['nodes_per_ray'], d['shadow']['prims_per_ray']))" >> $out 2>&1; }
run LJ_PRIM_THRESH=1 LJ_PRIM_THRESH_SHADOW=1
for t in 8 12 16 20 24; do run LJ_PRIM_THRESH=$t LJ_PRIM_THRESH_SHADOW=12; done
for t in 4 8 16 20; do run LJ_PRIM_THRESH=16 LJ_PRIM_THRESH_SHADOW=$t; done
for r in 12 16 26 30; do run LJ_REFILL_BELOW=$r LJ_REFILL_BELOW_SHADOW=$r; done
for g in 6 8 12 24 32; do run LJ_TRACE_GRID_MULT=$g; done
cat $out
