#!/bin/bash
# dev helper: time alternative builds of the library (build/libopal_*.so) on the bench workload
for so in build/libopal_*.so; do
  OPAL_B200_LIB=$PWD/$so python bench.py --nx ${NX:-200000} --steps 5 --warmup 2 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms']
print('$so', 'p-steps/s %.3e' % d['value'], 'push %.3f ms scatter %.3f ms' % (k['push'], k['scatter']), 'frac %.3f' % d['roofline']['frac'])"
done
