#!/bin/bash
O=gpurun_out/${1:-r2u}; mkdir -p $O
for NX in 708 1000 1416 2832; do
for lib in b2 b1; do
  MPSB_LIB_PATH=$PWD/build/exp/lib_$lib.so timeout 300 python tools/pcg_probe.py $NX > $O/probe_${lib}_nx$NX.log 2>&1
  echo "== $lib nx=$NX: $(grep 'us/it' $O/probe_${lib}_nx$NX.log | tail -1 | cut -c1-200)"
done; done
