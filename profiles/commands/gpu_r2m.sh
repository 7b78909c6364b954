#!/bin/bash
# stride of the all-to-all lines
O=gpurun_out/${1:-r2m}; mkdir -p $O
for NX in 1000 2832; do
 for st in 8 64 320 1056; do
  MPSB_GX_STRIDE=$st timeout 300 python tools/pcg_probe.py $NX > $O/probe_nx${NX}_st$st.log 2>&1
  echo "== nx=$NX stride=$st: $(grep 'us/it' $O/probe_nx${NX}_st$st.log | tail -1 | cut -c1-220)"
 done
done
