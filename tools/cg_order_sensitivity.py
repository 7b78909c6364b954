"""How much does the ITERATION COUNT of Jacobi-PCG at Eigen's stopping rule (|r|^2 < eps^2 |b|^2 on the recursively
updated residual) depend on nothing but the order of the floating-point sums? CPU only.

Takes the PPE the oracle assembles for scaled_dambreak(nx) (bit-identical to the reference's, tests/test_large_golden.py)
and runs the same recurrence (Eigen ConjugateGradient.h:28-91) in numpy under symmetric permutations of the unknowns
(identity = particle-id order as the reference; cell-major = the GPU's SORTED order; reverse; random). A permutation
changes only the order in which dot products and row sums are accumulated. Printed next to the reference's own count
(Eigen, serial left-to-right sums) and the solution difference between orderings.
    python tools/cg_order_sensitivity.py 400 > profiles/r2_cg_iteration_sensitivity_nx400.txt"""
import os
import sys

import numpy as np
import scipy.sparse as sp

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mps_b200 import cases as C
from oracle import refbind as R

nx = int(sys.argv[1]) if len(sys.argv) > 1 else 400
S, t, p, v = C.scaled_dambreak(nx)
orc = R.OracleMPS(S.to_c(), t, p, v)
for ph in (R.PH_NEIGHBORS, R.PH_GRAVITY, R.PH_VISCOSITY, R.PH_MOVE, R.PH_NEIGHBORS, R.PH_COLLISION, R.PH_NEIGHBORS,
           R.PH_BC, R.PH_SOURCE, R.PH_MATRIX, R.PH_DIRICHLET):
    orc.phase(ph)
start, col, val, rhs = orc.ppe()
n = len(t)
A = sp.csr_matrix((val, col, start), shape=(n, n))
x_ref, it_ref, err_ref, _ = orc.cg_solve()
print(f"nx={nx} n={n} nnz={len(col)}  reference order, Eigen's serial sums (oracle port == Eigen bit for bit): {it_ref} iterations")


def pcg(A, b):
    d = A.diagonal()
    inv = np.where(d != 0.0, 1.0 / np.where(d != 0.0, d, 1.0), 1.0)
    x = np.zeros_like(b)
    r = b.copy()
    rhs2 = float(b @ b)
    thr = max(np.finfo(float).eps ** 2 * rhs2, np.finfo(float).tiny)
    pz = inv * r
    absnew = float(r @ pz)
    pdir = pz.copy()
    it = 0
    while it < 2 * len(b):
        ap = A @ pdir
        alpha = absnew / float(pdir @ ap)
        x += alpha * pdir
        r -= alpha * ap
        if float(r @ r) < thr:
            break
        z = inv * r
        absold, absnew = absnew, float(r @ z)
        pdir = z + (absnew / absold) * pdir
        it += 1
    return x, it + 1 if float(r @ r) < thr else it


pos = p + 0.0
key = (np.floor((pos[:, 1] - S.domain_min[1]) / S.re_max).astype(np.int64) * 100000
       + np.floor((pos[:, 0] - S.domain_min[0]) / S.re_max).astype(np.int64))
orders = {"identity (numpy pairwise sums)": np.arange(n), "cell-major (the GPU's row order)": np.argsort(key, kind="stable"),
          "reverse": np.arange(n)[::-1].copy(), "random(1)": np.random.default_rng(1).permutation(n),
          "random(2)": np.random.default_rng(2).permutation(n)}
scale = np.abs(x_ref).max()
for name, perm in orders.items():
    Pm = sp.csr_matrix((np.ones(n), (np.arange(n), perm)), shape=(n, n))
    Ap = (Pm @ A @ Pm.T).tocsr()
    xp, it = pcg(Ap, rhs[perm])
    x = np.empty(n)
    x[perm] = xp
    print(f"  {name:34s}: {it:5d} iterations ({100.0 * (it - it_ref) / it_ref:+.1f} % vs reference)   "
          f"max|x - x_ref| / max|x_ref| = {np.abs(x - x_ref).max() / scale:.2e}")
