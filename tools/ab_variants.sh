#!/bin/bash
# A/B of fill-kernel build variants (build/lib_*.so): prints ms_per_step and the fill kernel's average launch time
for lib in "" build/lib_*.so; do
  HIST_B200_LIB=${lib:+$PWD/$lib} python bench.py --steps 10 --warmup 3 --cpu-pairs 1 --no-em 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('${lib:-default}', 'ms/step %.3f' % d['ms_per_step'], 'fill %.3f ms' % d['roofline']['avg_launch_ms'], 'GCUPS %.2f' % d['value'], 'e2e %.3f' % d['e2e']['value'])"
done
