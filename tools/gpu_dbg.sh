#!/bin/bash
O=gpurun_out/${1:-dbg}; mkdir -p $O
run() { tag=$1; shift; env "$@" timeout 120 python tools/pcg_probe.py ${NX:-200} > $O/probe_$tag.log 2>&1
  echo "== $tag rc=$?: $(grep 'us/it\|rror' $O/probe_$tag.log | tail -1 | cut -c1-250)"; }
run plain X=1
run blocking CUDA_LAUNCH_BLOCKING=1
run nol2 MPSB_L2_RESIDENT_KB=0
run stride8 MPSB_GX_STRIDE=8
run cap1 MPSB_CG_ITERATION_CAP=1
run cap3 MPSB_CG_ITERATION_CAP=3
NX=1000 run nx1000 X=1
