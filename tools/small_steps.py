import sys; sys.path.insert(0, ".")
from mps_b200 import capi, cases as C
S,t,p,v = C.load_case("dambreak_fine")
dev = capi.DeviceMPS(S,t,p,v); dev.load_initial_state()
for k in range(6): dev.step_forward(False)
dev.close()
