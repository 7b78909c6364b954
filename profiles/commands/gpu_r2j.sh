#!/bin/bash
# L2 residency, third pass: per-rank sizes of 4- and 2-GPU runs
O=gpurun_out/${1:-r2j}; mkdir -p $O
run() { tag=$1; shift; env "$@" timeout 300 python tools/pcg_probe.py $NX > $O/probe_$tag.log 2>&1
  echo "== nx=$NX $tag: $(grep 'us/it' $O/probe_$tag.log | tail -1 | cut -c1-220)"; }
for NX in 1416 2000; do
  run nx${NX}_off X=1
  for mb in 16 24 32 40; do run nx${NX}_keep$mb MPSB_L2_RESIDENT_MB=$mb; done
done
NX=1000
for mb in 16 24 40; do run nx${NX}_keep$mb MPSB_L2_RESIDENT_MB=$mb; done
