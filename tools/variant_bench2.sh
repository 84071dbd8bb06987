#!/bin/bash
# usage: variant_bench2.sh lib... ; runs each with classic and onesweep sort
for lib in "$@"; do
  for mode in classic onesweep; do
    GSR_SORT=$mode GSR_B200_LIB=$lib timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$lib','$mode', round(d['ms_per_step'],3), {k: round(v['ms'],3) for k,v in d['stages'].items() if 'sort' in k or 'scan' in k})"
  done
done
