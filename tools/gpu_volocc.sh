#!/bin/bash
set -u
timeout 600 python -m pytest tests/test_gpu_volpath.py -m gpu -x -q 2>&1 | tail -1
for rep in 1 2; do timeout 600 python tools/bench_scenes.py vol_cbox volpath_test3 volpath_test4 volpath_test6 vol_cbox_teapot hetvol hetvol_colored 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin)
print(' | '.join('%s %.1f' % (k, v['wavefront']['msamples_per_s']) for k,v in d.items()))"; done
