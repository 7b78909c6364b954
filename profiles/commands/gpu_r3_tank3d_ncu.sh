#!/bin/bash
# ncu --set full of the solve kernel on the 3-D tank at radius ratio 3.1 (265 852 particles), bounded to 60 iterations
O=gpurun_out/r3_tank3d; mkdir -p $O
MPSB_CG_ITERATION_CAP=60 timeout 600 ncu --set full --clock-control none --import-source on -k regex:pcg_kernel -c 1 -o $O/pcg_full_tank3d_r31 -f \
    python tools/tank3d_probe.py 3.1 1 > $O/ncu_full.log 2>&1; echo "ncu exit $?"
ncu -i $O/pcg_full_tank3d_r31.ncu-rep --page details --csv > $O/pcg_full_tank3d_r31_details.csv 2>/dev/null
python tools/ncu_summary.py $O/pcg_full_tank3d_r31.ncu-rep $O/pcg_full_tank3d_r31_summary.csv 2>&1 | tail -2
cut -c1-900 $O/pcg_full_tank3d_r31_summary.csv
