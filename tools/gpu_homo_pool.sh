#!/bin/bash
set -u
for s in vol_cbox volpath_test3 volpath_test4 volpath_test5 volpath_test6 vol_cbox_teapot volpath_test4_2; do echo "$(timeout 300 python tools/sweep_pool.py $s 64 1048576 2097152 4194304 | python -c "
import json,sys
d=json.load(sys.stdin); print(d['scene'], ' | '.join('%d Mi: %.1f Msamples/s (%d it)'%(int(k)>>20, v['msamples_per_s'], v['iterations']) for k,v in d['slots'].items()))")"; done
