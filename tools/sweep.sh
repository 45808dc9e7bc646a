#!/bin/bash
# dev helper: time alternative builds of the library (build/libopal_<tag>.so, see mkvariants.sh) on the bench workload
# usage: NX=1000000 tools/sweep.sh tag1 tag2 ...   (no tags: every build/libopal_*.so)
if [ $# -gt 0 ]; then libs=$(for t in "$@"; do echo build/libopal_$t.so; done); else libs=$(ls build/libopal_*.so); fi
for so in $libs; do
  OPAL_B200_LIB=$PWD/$so python bench.py --nx ${NX:-200000} --steps 5 --warmup 2 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms']
print('$so', 'p-steps/s %.3e' % d['value'], 'push %.3f ms scatter(x2) %.3f ms' % (k['push'], k['scatter']), 'frac %.3f' % d['roofline']['frac'])"
done
