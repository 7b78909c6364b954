#!/bin/bash
# L2 policy at the full single-GPU size (off by default there): evict_first only / 16 MB kept
O=gpurun_out/${1:-r2aa}; mkdir -p $O
run() { tag=$1; shift; env "$@" timeout 300 python tools/pcg_probe.py 2832 > $O/probe_$tag.log 2>&1
  echo "== $tag: $(grep 'us/it' $O/probe_$tag.log | tail -1 | cut -c1-200)"; }
run off X=1
run first_only MPSB_L2_MAX_ROWS=9000000 MPSB_L2_RESIDENT_KB=1
run keep16 MPSB_L2_MAX_ROWS=9000000 MPSB_L2_RESIDENT_KB=16384
run off2 X=1
