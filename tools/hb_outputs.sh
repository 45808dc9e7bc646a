#!/bin/bash
# dev helper: run two of the reference's example decks on the GPU and keep the energy files (+ a few histograms) under
# gpurun_out/deck_out/<deck>/ for a comparison with the CPU oracle driver's full run (tools/compare_deck_outputs.py)
for deck in ${DECKS:-hole_boring two_stream}; do
  rm -rf /tmp/deck_gpu; mkdir -p /tmp/deck_gpu && cp /root/repo/examples/$deck.yaml /tmp/deck_gpu/in.yaml
  (cd /tmp/deck_gpu && /root/repo/opal_b200/driver/opal_b200_driver in.yaml --quiet)
  out=/root/repo/gpurun_out/deck_out/$deck; mkdir -p $out
  cp /tmp/deck_gpu/*_energy.dat $out/
  for k in 10 20 30; do cp /tmp/deck_gpu/${k}_*_x-px.fits $out/ 2>/dev/null; done
  echo "$deck: $(ls $out | wc -l) files"
done
