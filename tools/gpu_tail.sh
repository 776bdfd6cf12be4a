#!/bin/bash
set -u
for t in 32768 65536 131072 262144; do for s in "sponza 8" "sponza 64" "cbox 4"; do
  LJ_WF_TAIL=$t timeout 300 python tools/profile_stages.py $s 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin)
print('tail %7s  %-8s %3d spp render %7.2f ms  iterations %d' % (sys.argv[1], d['scene'], d['spp'], d['render_ms_plain'], d['iterations']))" $t
done; done
