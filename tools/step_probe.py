"""A few steps of the scaled dam break with the solve capped (MPSB_CG_ITERATION_CAP), for ncu captures of the
NON-solve kernels at the bench size:
    MPSB_CG_ITERATION_CAP=8 ncu --set full -k regex:'neighbor|gradient|...' python tools/step_probe.py [nx] [steps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mps_b200 import capi, cases as C

nx = int(sys.argv[1]) if len(sys.argv) > 1 else 2832
S, t, p, v = C.scaled_dambreak(nx)
dev = capi.DeviceMPS(S, t, p, v)
dev.load_initial_state()
for k in range(int(sys.argv[2]) if len(sys.argv) > 2 else 2):
    dev.step_forward(k == 0)
    s = dev.last_stats
    print({k2: round(getattr(s, k2), 3) for k2 in ("ms_total", "ms_neighbors", "ms_explicit", "ms_collision", "ms_assemble",
                                                  "ms_solve", "ms_correct")}, "cg", s.cg_iterations, flush=True)
dev.close()
