"""Shared helpers for the parity tests."""
import glob
import os

import numpy as np

from mps_b200 import cases as C
from oracle import refbind as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
STATE_FIELDS = ["type", "position", "velocity", "acceleration", "pressure", "minimum_pressure",
                "number_density_ratio", "boundary_condition"]


def golden_files(case=None):
    pat = f"{case}_step*.npz" if case else "*_step*.npz"
    return sorted(glob.glob(os.path.join(GOLDEN, pat)))


def golden_case(path):
    return os.path.basename(path).rsplit("_step", 1)[0]


def load_golden(path):
    z = np.load(path)
    before = {n: z["before_" + n] for n in STATE_FIELDS}
    after = {n: z["after_" + n] for n in STATE_FIELDS}
    return z, before, after


def make_oracle(case, state=None):
    """OracleMPS (our restatement) for a shipped case, optionally forced to `state`."""
    S, t, p, v = C.load_case(case)
    o = R.OracleMPS(S.to_c(), t, p, v)
    if state is not None:
        o.set_state(state)
    return S, o


def non_ghost(state):
    return state["type"] != C.GHOST


def rel_err(a, b):
    """max |a-b| / max(|b|) — relative to the field's scale (per-step tolerance of north_star)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    scale = np.max(np.abs(b)) if b.size else 0.0
    if scale == 0.0:
        return float(np.max(np.abs(a - b))) if a.size else 0.0
    return float(np.max(np.abs(a - b)) / scale)


def sealed_box(nx=14, ny=14, l0=0.01):
    """A lattice box completely full of fluid inside a closed wall + dummy-wall shell: no particle is
    a free surface, so every Inner particle is Dirichlet-unconnected (reference src/mps.cpp:402-413)."""
    a = 3
    ix, iy = np.meshgrid(np.arange(-a, nx + a), np.arange(-a, ny + a), indexing="ij")
    ring = np.minimum(np.minimum(ix + a, nx + a - 1 - ix), np.minimum(iy + a, ny + a - 1 - iy))
    t = np.where(ring < 2, C.DUMMY_WALL, np.where(ring < 3, C.WALL, C.FLUID)).astype(np.int32).ravel()
    pos = np.zeros((t.size, 3))
    pos[:, 0] = ((ix + 0.5) * l0).ravel()
    pos[:, 1] = ((iy + 0.5) * l0).ravel()
    return t, pos, np.zeros_like(pos)


def elementwise_excess(a, b, tol, floor_frac=1e-3):
    """max over entries of |a-b| / (tol * max(|b_i|, floor_frac * max|b|)) — <= 1 means every entry agrees to
    `tol` RELATIVE TO ITSELF, except entries below floor_frac of the field's largest magnitude, which are held
    to the absolute bound tol * floor_frac * max|b| (a Krylov solve has a normwise, not a componentwise, error
    bound, so entries near zero cannot be asked for 1e-10 of themselves)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if b.size == 0:
        return 0.0
    scale = float(np.max(np.abs(b)))
    if scale == 0.0:
        return 0.0 if not np.any(a) else np.inf
    return float(np.max(np.abs(a - b) / (tol * np.maximum(np.abs(b), floor_frac * scale))))


def large_golden_files():
    return sorted(glob.glob(os.path.join(GOLDEN, "large_nx*.npz")),
                  key=lambda p: (int("".join(ch for ch in os.path.basename(p) if ch.isdigit())), p))


def sha(*arrays):
    import hashlib

    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def large_case(z):
    """Regenerates the inputs of a tests/golden/large_nx*.npz fixture (tests/golden/make_large_golden.py)."""
    S, t, p, v = C.scaled_dambreak(int(z["nx"]))
    if int(z["jitter"]):
        p, v = C.jittered(S, t, p, seed=int(z["nx"]))
    return S, t, p, v
