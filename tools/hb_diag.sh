#!/bin/bash
# dev helper: where hole_boring.yaml spends its time -- per-interval step time (OPAL_STATS=1), per-kernel device time (OPAL_STATS=2,
# step by step)
mkdir -p /tmp/hb && cp examples/hole_boring.yaml /tmp/hb/in.yaml
cd /tmp/hb
for mode in "OPAL_STATS=2" "OPAL_STATS=1"; do
  echo "== $mode"
  env $mode /root/repo/opal_b200/driver/opal_b200_driver in.yaml --quiet 2>&1 | grep -v "^Output"
done
