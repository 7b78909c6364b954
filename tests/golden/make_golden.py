"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libmps_ref.so, built by
`make -C oracle ref` from /root/reference). Each fixture is a teacher-forcing pair: the reference's
full state before a step, and everything observable after it. Run from the repo root:
    python tests/golden/make_golden.py
The fixtures are committed; the GPU box (which has no /root/reference) reads only them.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from mps_b200 import cases as C  # noqa: E402
from oracle import refbind as R  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

# case -> steps at which a (before, after) pair is dumped
PLAN = {
    "dambreak": [0, 1, 150, 324, 430, 700],       # 324/430: steps in which the reference ghosts a particle
    "hydrostatic": [0, 30],
    "dambreak_fine": [0, 60],
    "dambreak_fine_original": [0, 20],
    "dambreak_fine_higher_order": [0, 20],
    "tank3d": [0, 5],                              # generated 3-D case (mps_b200.cases.tank3d); the reference ships none
    "tank3d_r31": [0, 5],                          # the same tank at the reference's shipped radius ratio 3.1 (~120 neighbours per particle)
}


def nb_digest(start, ids, dist):
    h = hashlib.sha256()
    h.update(start.tobytes()); h.update(ids.tobytes()); h.update(dist.tobytes())
    return h.hexdigest()


def main():
    only = sys.argv[1:]  # optional: regenerate only these cases
    for case, steps in PLAN.items():
        if only and case not in only:
            continue
        if case.startswith("tank3d"):
            import tempfile

            S, t, p, v = C.load_case(case)
            yml = C.write_settings_yaml(S, os.path.join(tempfile.mkdtemp(prefix="mpsb_golden_"), "settings.yml"))
            ref = R.RefMPS(yml, arrays=(t, p, v))
        else:
            yml, _ = C.case_paths(case)
            S = C.load_settings(yml)
            ref = R.RefMPS(yml, os.path.join(ROOT, S.input_csv_path))
        k = 0
        for target in steps:
            while k < target:
                ref.step(False)
                k += 1
            before = ref.get_state()
            dbg0 = ref.debug_log_count
            # phase by phase so the PPE and the iteration count can be recorded mid-step
            for ph in (R.PH_NEIGHBORS, R.PH_GRAVITY, R.PH_VISCOSITY, R.PH_MOVE, R.PH_NEIGHBORS, R.PH_COLLISION,
                       R.PH_NEIGHBORS):
                ref.phase(ph)
            nstart, nids, ndist = ref.neighbors()
            for ph in (R.PH_BC, R.PH_SOURCE, R.PH_MATRIX, R.PH_DIRICHLET):
                ref.phase(ph)
            pstart, pcol, pval, prhs = ref.ppe()
            _, iters, err, info = ref.cg_solve()
            for ph in (R.PH_SOLVE, R.PH_CLAMP, R.PH_MINP, R.PH_GRADIENT, R.PH_MOVE2):
                ref.phase(ph)
            export = (k % 2 == 0)
            if export or S.higher_order_source_on:
                ref.phase(R.PH_NUMBER_DENSITY)
            courant = ref.phase(R.PH_COURANT)
            k += 1
            after = ref.get_state()
            out = {("before_" + n): a for n, a in before.items()}
            out.update({("after_" + n): a for n, a in after.items()})
            out.update(dict(
                step=k - 1, export=int(export), courant=courant, cg_iterations=iters, cg_error=err,
                nnz=len(pcol), n_inner=int((after["boundary_condition"] == 2).sum()),
                n_surface=int((after["boundary_condition"] == 1).sum()),
                debug_events=ref.debug_log_count - dbg0,
                nbr_total=len(nids), nbr_digest=nb_digest(nstart, nids, ndist),
                ppe_row_start=pstart, ppe_col=pcol, ppe_val=pval, ppe_rhs=prhs,
            ))
            path = os.path.join(HERE, f"{case}_step{k - 1:04d}.npz")
            np.savez_compressed(path, **out)
            print(case, k - 1, "cg", iters, "nnz", len(pcol), "dbg", ref.debug_log_count - dbg0,
                  "%.0f kB" % (os.path.getsize(path) / 1024))


if __name__ == "__main__":
    main()
