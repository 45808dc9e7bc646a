#!/bin/bash
# dev helper: run hole_boring.yaml on the GPU and keep the energy files (+ a few histograms) for a comparison with the CPU oracle's run
mkdir -p /tmp/hb_gpu && cp examples/hole_boring.yaml /tmp/hb_gpu/in.yaml && cd /tmp/hb_gpu && /root/repo/opal_b200/driver/opal_b200_driver in.yaml --quiet
mkdir -p /root/repo/gpurun_out/hb_out && cp /tmp/hb_gpu/*_energy.dat /root/repo/gpurun_out/hb_out/ && cp /tmp/hb_gpu/10_electron_x-px.fits /tmp/hb_gpu/20_electron_x-px.fits /tmp/hb_gpu/30_electron_x-px.fits /tmp/hb_gpu/30_carbon_x-px.fits /tmp/hb_gpu/20_grid.dat /root/repo/gpurun_out/hb_out/
ls /root/repo/gpurun_out/hb_out | wc -l
