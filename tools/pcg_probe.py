"""Diagnostic: time the PCG solve alone on the scaled dam break and print the per-sweep cycle split.
    python tools/pcg_probe.py [nx] [iteration cap via MPSB_CG_ITERATION_CAP]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mps_b200 import capi, cases as C

nx = int(sys.argv[1]) if len(sys.argv) > 1 else 2832
S, t, p, v = C.scaled_dambreak(nx)
dev = capi.DeviceMPS(S, t, p, v)
dev.load_initial_state()
for ph in (capi.PH_NEIGHBORS, capi.PH_EXPLICIT, capi.PH_NEIGHBORS, capi.PH_COLLISION, capi.PH_NEIGHBORS,
           capi.PH_FLAGS_SOURCE, capi.PH_MATRIX, capi.PH_DIRICHLET):
    st = dev.run_phase(ph)
for rep in range(3):
    t0 = time.perf_counter()
    st = dev.run_phase(capi.PH_SOLVE)
    dt = time.perf_counter() - t0
    cyc = dev.cg_cycles()
    tot = sum(cyc) or 1
    it = max(st.cg_iterations, 1)
    print(f"nx={nx} n_inner={st.n_inner} nnz={st.nnz} it={st.cg_iterations} {1e6*dt/it:.1f} us/it  "
          f"A {100*cyc[0]/tot:.1f}% B {100*cyc[1]/tot:.1f}% C {100*cyc[2]/tot:.1f}% sync {100*cyc[3]/tot:.1f}%  "
          f"cycles/it A {cyc[0]/it:.0f} B {cyc[1]/it:.0f} C {cyc[2]/it:.0f} sync {cyc[3]/it:.0f}", flush=True)
b = dev.cg_block_cycles()
import numpy as np
it = max(st.cg_iterations, 1)
for k, name in enumerate(("A", "B", "C", "sync")):
    v = b[k] / it
    print(f"per-CTA cycles/it {name}: min {v.min():.0f} mean {v.mean():.0f} max {v.max():.0f}")
order = np.argsort(b[0])
print("slowest A CTAs (block, smid, cycles/it):", [(int(i), int(b[4][i]), int(b[0][i] / it)) for i in order[-8:]])
print("fastest A CTAs (block, smid, cycles/it):", [(int(i), int(b[4][i]), int(b[0][i] / it)) for i in order[:8]])
