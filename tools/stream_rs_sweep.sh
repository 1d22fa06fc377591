#!/bin/bash
# A/B: streaming NConv kernel at several rows-per-unit settings (forward only).
for rs in 0 8 12 17 24 32 48 64; do
  echo "RS=$rs"
  NCONV_B200_STREAM_RS=$rs python tools/nconv2d_sweep.py --no-bwd 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l)
    if 'k' in d and d['ci']*d['co']<=16 and d['co']<8: print('  ',d['k'],d['ci'],d['co'],d['fwd_ms'],d['fwd_hbm_frac'])
"
done
