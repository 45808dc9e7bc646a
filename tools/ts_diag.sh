#!/bin/bash
# dev helper: time loop of the two small example decks with OPAL_STATS=1 (launches, bursts, halts, re-layouts, host seconds)
for deck in two_stream hole_boring; do
mkdir -p /tmp/$deck && cp examples/$deck.yaml /tmp/$deck/in.yaml && cd /tmp/$deck
for hm in ${HEAVY_MINS:-1024}; do
echo "== $deck OPAL_HEAVY_MIN=$hm"; OPAL_HEAVY_MIN=$hm OPAL_STATS=1 /root/repo/opal_b200/driver/opal_b200_driver in.yaml 2>&1 | grep "stats\|time loop" | cut -c1-200
done
cd /root/repo
done
