#!/bin/bash
# per-kernel time breakdown of a deck's time loop (ncu launch list; cold-cache, serialised: compare SHARES): tools/profile_deck.sh <deck> <steps> <tag>
set -e
D=/tmp/opal_prof_$3; mkdir -p $D; cp examples/$1.yaml $D/in.yaml
opal_b200/driver/opal_b200_driver $D/in.yaml --outputs 1 --steps $2 --quiet > gpurun_out/$3_plain.log 2>&1
OPAL_BURST=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/$3_launches.csv opal_b200/driver/opal_b200_driver $D/in.yaml --outputs 1 --steps $2 --quiet > gpurun_out/$3_ncu.log 2>&1
python - <<PY
import csv,collections
rows=[r for r in csv.reader(open("gpurun_out/$3_launches.csv")) if len(r)>14 and r[0].isdigit()]
agg=collections.OrderedDict(); tot=0.0
for r in rows:
    name=r[4].split("(")[0].replace("void ","").replace("opal::fast::","").replace("opal::","")
    a=agg.setdefault(name,[0,0.0]); a[0]+=1; a[1]+=float(r[14]); tot+=float(r[14])
print("$1: %d launches, %.1f ms of kernel time" % (len(rows), tot/1e6))
for k,(n,t) in sorted(agg.items(), key=lambda kv:-kv[1][1]): print("  %-44s n=%5d  %8.2f ms  %5.1f %%  avg %8.2f us" % (k[:44], n, t/1e6, 100*t/tot, t/n/1e3))
PY
