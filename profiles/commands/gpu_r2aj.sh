#!/bin/bash
# ring geometry at the per-rank sizes of multi-GPU runs (round 1 tuned it at 4 M rows only)
O=gpurun_out/${1:-r2aj}; mkdir -p $O
for NX in 1000 1416; do for lib in build/exp/lib_*.so; do
  name=$(basename $lib .so)
  MPSB_LIB_PATH=$PWD/$lib timeout 300 python tools/pcg_probe.py $NX > $O/${name}_nx$NX.log 2>&1
  echo "== $name nx=$NX: $(grep 'us/it' $O/${name}_nx$NX.log | tail -1 | cut -c1-200)"
done; done
