#!/bin/bash
# dev helper: time alternative builds of the library (build/libopal_<tag>.so, see mkvariants.sh) on the bench workload
# usage: NX=1000000 tools/sweep.sh tag1 tag2 ...   (no tags: every build/libopal_*.so)
if [ $# -gt 0 ]; then libs=$(for t in "$@"; do echo build/libopal_$t.so; done); else libs=$(ls build/libopal_*.so); fi
for so in $libs; do
  OPAL_B200_LIB=$PWD/$so python bench.py --nx ${NX:-1000000} --steps 10 --warmup 3 --reps 3 --no-cpu-baseline --no-parity --no-config5 2>&1 | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms']
print('$so', 'p-steps/s %.3e' % d['value'], 'ms/step %.3f' % d['ms_per_step'], 'push %.3f ms scatter %.3f ms' % (k['push'], k['scatter']), 'frac %.3f' % d['roofline']['frac'], 'clk', d['clocks']['sm_mhz'], d['clocks']['power_w_median'])"
done
