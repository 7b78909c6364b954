"""A few steps of a shipped case, for an ncu launch list of the small-case step:
    MPSB_NO_STEP_GRAPH=1 ncu --metrics gpu__time_duration.sum --clock-control none --csv python tools/small_steps.py [case] [steps]"""
import sys
sys.path.insert(0, ".")
from mps_b200 import capi, cases as C

case = sys.argv[1] if len(sys.argv) > 1 else "dambreak_fine"
S, t, p, v = C.load_case(case)
dev = capi.DeviceMPS(S, t, p, v)
dev.load_initial_state()
for k in range(int(sys.argv[2]) if len(sys.argv) > 2 else 6):
    dev.step_forward(False)
dev.close()
