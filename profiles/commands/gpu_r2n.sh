#!/bin/bash
O=gpurun_out/${1:-r2n}; mkdir -p $O
run() { tag=$1; shift; env "$@" timeout 300 python tools/pcg_probe.py ${NX:-1000} > $O/probe_$tag.log 2>&1
  echo "== $tag: $(grep 'us/it' $O/probe_$tag.log | tail -1 | cut -c1-220)"; grep "per-CTA cycles/it sync" $O/probe_$tag.log; }
run base MPSB_GX_STRIDE=320
run s8 MPSB_GX_STRIDE=8
NX=2832 run nx2832 MPSB_GX_STRIDE=320
