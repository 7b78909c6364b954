"""Second pin for the ITERATION COUNT of the large-N solves: tests/golden/large_pairwise_iterations.json.

Eigen's CG sums its dot products serially; at 1e5-4e6 unknowns that rounding alone delays the crossing of the
eps threshold by several per cent against the same recurrence with tree/pairwise sums (tools/cg_order_sensitivity.py,
profiles/r2_cg_iteration_sensitivity_nx400.txt: 999 vs 936 at nx = 400 under every ordering). The GPU sums in trees, so
its count is compared (a) with Eigen's within a stated band and (b) with THIS count — the reference's own PPE (digest
equal to the compiled reference's, checked below) solved on the CPU by the identical recurrence
(ConjugateGradient.h:28-91) with numpy's pairwise sums — within 1 %.
    python tests/golden/make_pairwise_iterations.py 400 400j 800 800j 1416 [2832]"""
import json
import os
import sys

import numpy as np
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import refbind as R  # noqa: E402
from tests import util as U  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "large_pairwise_iterations.json")


def pcg_iterations(A, b):
    d = A.diagonal()
    inv = np.where(d != 0.0, 1.0 / np.where(d != 0.0, d, 1.0), 1.0)
    x = np.zeros_like(b)
    r = b.copy()
    thr = max(np.finfo(float).eps ** 2 * float(b @ b), np.finfo(float).tiny)
    z = inv * r
    absnew = float(r @ z)
    pdir = z.copy()
    it = 0
    while it < 2 * len(b):
        ap = A @ pdir
        alpha = absnew / float(pdir @ ap)
        x += alpha * pdir
        r -= alpha * ap
        it += 1
        if float(r @ r) < thr:
            break
        z = inv * r
        absold, absnew = absnew, float(r @ z)
        pdir = z + (absnew / absold) * pdir
    return it, x


def main():
    out = json.load(open(OUT)) if os.path.exists(OUT) else {}
    for tag in sys.argv[1:]:
        z = np.load(os.path.join(ROOT, "tests", "golden", f"large_nx{tag}.npz"))
        S, t, p, v = U.large_case(z)
        orc = R.OracleMPS(S.to_c(), t, p, v)
        for ph in (R.PH_NEIGHBORS, R.PH_GRAVITY, R.PH_VISCOSITY, R.PH_MOVE, R.PH_NEIGHBORS, R.PH_COLLISION, R.PH_NEIGHBORS,
                   R.PH_BC, R.PH_SOURCE, R.PH_MATRIX, R.PH_DIRICHLET):
            orc.phase(ph)
        ppe = orc.ppe()
        assert U.sha(*ppe) == str(z["ppe_digest"]), "the oracle's PPE is not the reference's"
        start, col, val, rhs = ppe
        n = len(t)
        it, x = pcg_iterations(sp.csr_matrix((val, col, start), shape=(n, n)), rhs)
        ids = z["sample_ids"]
        out[tag] = {"cg_iterations_pairwise": it, "cg_iterations_eigen": int(z["cg_iterations"]),
                    "sample_max_rel_vs_eigen": float(np.abs(x[ids] - z["sample_solve"]).max() / float(z["p_norminf"]))}
        print(tag, out[tag], flush=True)
        json.dump(out, open(OUT, "w"), indent=1, sort_keys=True)
        orc.close()


if __name__ == "__main__":
    main()
