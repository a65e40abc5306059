#!/bin/bash
# Tuning aid: throughput of mcb_run_steps for every (warps per chain, lanes per step) variant of k_run_spec.
for t in "1,32" "1,16" "2,16" "1,8" "2,8"; do
  echo -n "W,GL=$t : "
  MCB200_TUNE=$t timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('%.4g site-pairs/s  %.4g moves/s  frac %.3f  e2e %.4g  %.2f ms/launch' % (d['value'], d['trial_moves_per_s'], d['roofline']['frac'], d['e2e']['value'], d['ms_per_step']))" 2>&1 | tail -1
done
