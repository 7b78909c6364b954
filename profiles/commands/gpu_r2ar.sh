#!/bin/bash
O=gpurun_out/${1:-r2ar}; mkdir -p $O
for lib in unroll2 unroll4; do
  MPSB_CG_SMALL_ROWS=9000000 MPSB_LIB_PATH=$PWD/build/exp/lib_$lib.so timeout 200 python tools/pcg_probe.py 2832 > $O/probe_$lib.log 2>&1
  echo "== $lib (16x16 ring forced, 4 M rows): $(grep 'us/it' $O/probe_$lib.log | tail -1 | cut -c1-200)"
done
