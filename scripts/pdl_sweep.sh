#!/bin/bash
# A/B the programmatic-dependent-launch mask: one bench run per mask, device-resident value only
for m in 0x0 0x3ff 0x1ff 0x00e 0x1c0 0x004 0x002 0x100; do
  PDT_PDL=$m python bench.py --no-extras --steps 200 --warmup 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('mask $m  us/step %.1f  e2e %.1f' % (1e3*d['ms_per_step'], 1e3*d['e2e']['ms_per_step']))"
done
