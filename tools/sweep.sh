#!/bin/bash
# Variant sweep of the headline config on one B200 (bench value only).
for k in ldg128 ldg256; do for t in 128 256 512; do for c in 0 1 2 4; do for u in 1 2; do
  echo -n "$k t=$t c=$c u=$u : "
  python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu --kernel $k --threads $t --ctas $c --unroll $u 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(d['ms_per_step'], d['roofline']['achieved'], d['roofline']['frac'])
"
done; done; done; done
