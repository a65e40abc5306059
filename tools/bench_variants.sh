#!/bin/bash
# tuning aid: bench.py's headline leg for every montecarlo_b200/libv_*.so (variant builds), one line each
for so in montecarlo_b200/libv_*.so; do
  MCB200_LIB=$so python bench.py --steps 8 --warmup 3 --no-cpu-baseline --target-chains ${TARGET_CHAINS:-0} 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
t=d.get('target_box')
print('$so', '%.4g'%d['value'], '%.4f'%d['roofline']['frac'], d['oracle_check']['ok'], ('target %.4g %s'%(t['value'], t['oracle_check']['ok'])) if t else '')"
done
