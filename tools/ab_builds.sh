#!/bin/bash
# A/B of builds with different compile-time parameters (SPHB_DEFS=... python -m libclsph_b200.build):
# per-pass times of tools/ab_neighbors.py for every libclsph_b200/build/defs_*/libsphb.so
for d in libclsph_b200/build/defs_*; do echo "== $d"; SPHB_LIB=$d/libsphb.so timeout 120 python tools/ab_neighbors.py --steps 5 ${AB_ARGS:---variants 0} 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('variant %d density %.3f forces %.3f step %.3f'%(d['variant'],d['passes']['density'],d['passes']['forces'],d['ms_per_step']))
"; done
