"""3-D solve probe: µs per CG iteration of the generated open tank at a radius ratio, under the library's
environment switches as set by the caller.   python tools/tank3d_probe.py [ratio] [steps]"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    from mps_b200 import capi, cases as C

    ratio = float(sys.argv[1]) if len(sys.argv) > 1 else 3.1
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    geo = ((64, 64, 40), (128, 80, 40)) if ratio > 2.5 else ((32, 40, 24), (64, 56, 24))
    S, t, p, v = C.tank3d(n_fluid=geo[0], n_tank=geo[1], ratio=ratio, jitter=0.05)
    dev = capi.DeviceMPS(S, t, p, v)
    dev.load_initial_state()
    for _ in range(3):
        dev.step_forward(False)
    its, solve, tot = [], [], []
    for _ in range(steps):
        dev.step_forward(False)
        st = dev.last_stats
        its.append(st.cg_iterations); solve.append(st.ms_solve); tot.append(st.ms_total)
    st = dev.last_stats
    cyc = dev.cg_cycles()
    b = 12.0 * st.nnz + 108.0 * st.n_inner
    us = 1e3 * float(np.sum(solve)) / max(1, int(np.sum(its)))
    print(json.dumps({"env": {k: v for k, v in os.environ.items() if k.startswith("MPSB_")}, "ratio": ratio, "particles": len(t),
                      "n_inner": st.n_inner, "nnz": st.nnz, "max_neighbors": st.max_neighbors, "us_per_cg_iteration": us,
                      "algorithmic_gbs": b / us / 1e3, "ms_total": float(np.mean(tot)), "ms_solve": float(np.mean(solve)),
                      "cycle_split_ABCsync": [c / max(1, sum(cyc)) for c in cyc]}))
    dev.close()


if __name__ == "__main__":
    main()
