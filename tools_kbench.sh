#!/bin/bash
# quick kernel iteration benchmark: env passes through
python bench.py --steps 5 --warmup 2 --no-cpu-baseline --no-e2e 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('ms/step %.2f  avg_rhs_ms %.3f  kfrac %.3f  stepfrac %.3f' % (d['ms_per_step'], d['roofline']['kernel_ms_per_substep'], d['roofline']['frac'], d['roofline']['step_frac']))
    else: print(l.strip())
"
