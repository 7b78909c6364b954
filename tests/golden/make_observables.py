"""Generates tests/golden/observables.json from the UNMODIFIED reference (oracle/_ref/libmps_ref.so):
free-running time series of the observables BASELINE.json names — the dam-break front position and the
hydrostatic bottom-row mean pressure. Run from the repo root:  python tests/golden/make_observables.py
Definitions (also used by tests/test_gpu_parity.py):
  x_front   = max x over Fluid particles
  p_bottom  = mean pressure over Fluid particles with y < l0 (the lowest fluid row of the lattice at t = 0)
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from mps_b200 import cases as C  # noqa: E402
from oracle import refbind as R  # noqa: E402

PLAN = {  # case -> steps after which the observables are recorded
    "dambreak": [50, 100, 150, 200, 250],
    "dambreak_fine": [100, 200, 400, 600],
    "hydrostatic": [10, 50, 100, 250],
}


def observe(S, st):
    fluid = st["type"] == C.FLUID
    bottom = fluid & (st["position"][:, 1] < S.particle_distance)
    return {"x_front": float(st["position"][fluid, 0].max()), "y_top": float(st["position"][fluid, 1].max()),
            "p_bottom": float(st["pressure"][bottom].mean()) if bottom.any() else 0.0,
            "n_fluid": int(fluid.sum())}


def main():
    out = {}
    for case, marks in PLAN.items():
        yml, _ = C.case_paths(case)
        S = C.load_settings(yml)
        ref = R.RefMPS(yml, os.path.join(ROOT, S.input_csv_path))
        series, k = {}, 0
        for m in marks:
            while k < m:
                ref.step(False)
                k += 1
            series[str(m)] = observe(S, ref.get_state(["type", "position", "pressure"]))
            print(case, m, series[str(m)], flush=True)
        out[case] = {"dt": S.dt, "l0": S.particle_distance, "series": series}
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "observables.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
