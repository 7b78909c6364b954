#!/bin/bash
# ncu --set full of the solve kernel at the bench size (bounded to 300 iterations so that ~40 replays stay short)
O=gpurun_out/${1:-r2ac}; mkdir -p $O
MPSB_CG_ITERATION_CAP=300 timeout 900 ncu --set full --clock-control none --import-source on -k regex:pcg_kernel -c 1 -o $O/pcg_full_nx2832_cap300 -f \
    python tools/pcg_probe.py 2832 > $O/ncu_full.log 2>&1; echo "ncu exit $?"
ls -la $O | tail -3
ncu -i $O/pcg_full_nx2832_cap300.ncu-rep --page details --csv > $O/pcg_full_nx2832_cap300_details.csv 2>/dev/null
python tools/ncu_summary.py $O/pcg_full_nx2832_cap300.ncu-rep $O/pcg_full_nx2832_cap300_summary.csv 2>&1 | tail -2
cat $O/pcg_full_nx2832_cap300_summary.csv | cut -c1-600
