#!/bin/bash
# dev helper: the two small example decks with OPAL_STATS=1 (time loop, outputs, launches, bursts, halts, re-layouts, host seconds)
# HEAVY_MINS="0 4096": one run per value of OPAL_HEAVY_MIN
for deck in two_stream hole_boring; do
mkdir -p /tmp/$deck && cp examples/$deck.yaml /tmp/$deck/in.yaml && cd /tmp/$deck
for hm in ${HEAVY_MINS:-4096}; do
echo "== $deck OPAL_HEAVY_MIN=$hm"; OPAL_HEAVY_MIN=$hm OPAL_STATS=1 /root/repo/opal_b200/driver/opal_b200_driver in.yaml 2>&1 | grep "stats\|time loop\|outputs:" | cut -c1-260
done
cd /root/repo
done
