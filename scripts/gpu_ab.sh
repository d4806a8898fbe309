#!/bin/bash
# A/B of several builds of the library (distributed-neural-networks_b200/ab_*.so) on the same box:
# part-0 kernel time from the bench line, twice each.  The last one listed stays installed.
cd distributed-neural-networks_b200
cp libb200pipe.so /tmp/keep.so
for round in 1 2; do for f in ab_*.so; do
  cp $f libb200pipe.so
  (cd .. && timeout 300 python bench.py --steps 300 --warmup 10 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$f', round(d['roofline']['us_per_launch'],2), 'us part0;', round(d['value']/1e6,2), 'M samples/s;', d['e2e']['max_abs_diff_vs_oracle'])")
done; done
cp /tmp/keep.so libb200pipe.so
