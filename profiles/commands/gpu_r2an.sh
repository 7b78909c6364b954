#!/bin/bash
O=gpurun_out/${1:-r2an}; mkdir -p $O
run() { tag=$1; shift; env "$@" timeout 300 python tools/pcg_probe.py ${NX:-2832} > $O/probe_$tag.log 2>&1
  echo "== $tag: $(grep 'us/it' $O/probe_$tag.log | tail -1 | cut -c1-200)"; }
run v0 X=1
run v1 MPSB_CG_SMALL_ROWS=9000000
run v1_l2 MPSB_CG_SMALL_ROWS=9000000 MPSB_L2_MAX_ROWS=9000000
run v0_again X=1
NX=2000 run nx2000_v0 MPSB_CG_SMALL_ROWS=0
NX=2000 run nx2000_v1 MPSB_CG_SMALL_ROWS=9000000
