"""ctypes bindings for the two CPU checkers. TEST INFRASTRUCTURE — only tests/, bench.py's
cpu_baseline / --impl reference legs and __graft_entry__.smoke() may import this.

  RefMPS    : oracle/_ref/libmps_ref.so  — the UNMODIFIED reference compiled from /root/reference
              (oracle/ref_harness.cpp + oracle/Makefile).
  OracleMPS : oracle/libmps_oracle.so    — our restatement (oracle/mps_oracle.cpp).
Both expose the same Python surface so a parity test can be run against either.
"""
from __future__ import annotations

import ctypes as ct
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(HERE, "_ref", "libmps_ref.so")
REF_EXE = os.path.join(HERE, "_ref", "main.exe")
ORACLE_LIB = os.path.join(HERE, "libmps_oracle.so")

import sys  # noqa: E402

sys.path.insert(0, os.path.dirname(HERE))
from mps_b200.cases import CSettings  # noqa: E402


class CFields(ct.Structure):
    _fields_ = [
        ("type", ct.c_void_p),
        ("position", ct.c_void_p),
        ("velocity", ct.c_void_p),
        ("acceleration", ct.c_void_p),
        ("pressure", ct.c_void_p),
        ("minimum_pressure", ct.c_void_p),
        ("number_density_ratio", ct.c_void_p),
        ("boundary_condition", ct.c_void_p),
    ]


FIELD_SPECS = [
    ("type", np.int32, 1),
    ("position", np.float64, 3),
    ("velocity", np.float64, 3),
    ("acceleration", np.float64, 3),
    ("pressure", np.float64, 1),
    ("minimum_pressure", np.float64, 1),
    ("number_density_ratio", np.float64, 1),
    ("boundary_condition", np.int32, 1),
]


def alloc_fields(n, names=None):
    """dict of numpy arrays + the CFields struct pointing at them."""
    arrs, cf = {}, CFields()
    for name, dt, w in FIELD_SPECS:
        if names is not None and name not in names:
            continue
        a = np.zeros((n, w) if w > 1 else n, dtype=dt)
        arrs[name] = a
        setattr(cf, name, a.ctypes.data)
    return arrs, cf


def fields_from(arrs):
    cf = CFields()
    keep = {}
    for name, dt, w in FIELD_SPECS:
        if name in arrs and arrs[name] is not None:
            a = np.ascontiguousarray(arrs[name], dtype=dt)
            keep[name] = a
            setattr(cf, name, a.ctypes.data)
    return keep, cf


def have_ref():
    return os.path.exists(REF_LIB)


def have_oracle():
    return os.path.exists(ORACLE_LIB)


# reference phase ids (oracle/ref_harness.cpp ref_phase)
(PH_NEIGHBORS, PH_GRAVITY, PH_VISCOSITY, PH_MOVE, PH_COLLISION, PH_BC, PH_SOURCE, PH_MATRIX, PH_DIRICHLET,
 PH_SOLVE, PH_CLAMP, PH_MINP, PH_GRADIENT, PH_MOVE2, PH_NUMBER_DENSITY, PH_COURANT) = range(16)


class _Base:
    prefix = ""
    lib = None

    def _f(self, name, restype=None, argtypes=None):
        fn = getattr(self.lib, self.prefix + name)
        fn.restype = restype
        if argtypes is not None:
            fn.argtypes = argtypes
        return fn

    def _bind(self):
        vp = ct.c_void_p
        self._destroy = self._f("destroy", None, [vp])
        self._n = self._f("num_particles", ct.c_int64, [vp])
        self._get_settings = self._f("get_settings", None, [vp, ct.POINTER(CSettings)])
        self._get_state = self._f("get_state", None, [vp, ct.POINTER(CFields)])
        self._set_state = self._f("set_state", None, [vp, ct.POINTER(CFields)])
        self._phase = self._f("phase", ct.c_double, [vp, ct.c_int])
        self._step = self._f("step", ct.c_double, [vp, ct.c_int])
        self._get_neighbors = self._f("get_neighbors", ct.c_int64, [vp, vp, vp, vp, ct.c_int64])
        self._get_ppe = self._f("get_ppe", ct.c_int64, [vp, vp, vp, vp, vp, ct.c_int64])
        self._cg_solve = self._f("cg_solve", ct.c_int, [vp, vp, ct.POINTER(ct.c_double), ct.POINTER(ct.c_int)])
        self._get_conn = self._f("get_dirichlet_connected", None, [vp, vp])
        self._set_popt = self._f("set_pressure_options", None,
                                 [vp, ct.c_int, ct.c_double, ct.c_int, ct.c_double, ct.c_double])
        self._dbg = self._f("debug_log_count", ct.c_int, [vp])

    def close(self):
        if getattr(self, "h", None):
            self._destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def n(self):
        return int(self._n(self.h))

    @property
    def debug_log_count(self):
        return int(self._dbg(self.h))

    def settings(self) -> CSettings:
        s = CSettings()
        self._get_settings(self.h, ct.byref(s))
        return s

    def set_pressure_options(self, qc_on, alpha, ho_on, gamma, relax):
        self._set_popt(self.h, int(qc_on), float(alpha), int(ho_on), float(gamma), float(relax))

    def get_state(self, names=None):
        arrs, cf = alloc_fields(self.n, names)
        self._get_state(self.h, ct.byref(cf))
        return arrs

    def set_state(self, arrs):
        keep, cf = fields_from(arrs)
        self._set_state(self.h, ct.byref(cf))

    def phase(self, ph):
        return float(self._phase(self.h, int(ph)))

    def step(self, is_time_to_export=False):
        return float(self._step(self.h, int(bool(is_time_to_export))))

    def neighbors(self):
        n = self.n
        start = np.zeros(n + 1, dtype=np.int64)
        total = self._get_neighbors(self.h, start.ctypes.data, None, None, 0)
        ids = np.zeros(total, dtype=np.int32)
        dist = np.zeros(total, dtype=np.float64)
        self._get_neighbors(self.h, start.ctypes.data, ids.ctypes.data, dist.ctypes.data, total)
        return start, ids, dist

    def ppe(self):
        n = self.n
        start = np.zeros(n + 1, dtype=np.int64)
        rhs = np.zeros(n, dtype=np.float64)
        nnz = self._get_ppe(self.h, start.ctypes.data, None, None, rhs.ctypes.data, 0)
        col = np.zeros(nnz, dtype=np.int32)
        val = np.zeros(nnz, dtype=np.float64)
        self._get_ppe(self.h, start.ctypes.data, col.ctypes.data, val.ctypes.data, rhs.ctypes.data, nnz)
        return start, col, val, rhs

    def cg_solve(self):
        x = np.zeros(self.n)
        err = ct.c_double()
        info = ct.c_int()
        it = self._cg_solve(self.h, x.ctypes.data, ct.byref(err), ct.byref(info))
        return x, int(it), float(err.value), int(info.value)

    def dirichlet_connected(self):
        out = np.zeros(self.n, dtype=np.int32)
        self._get_conn(self.h, out.ctypes.data)
        return out


class RefMPS(_Base):
    """The compiled reference. Build with `make -C oracle ref` (needs /root/reference)."""

    prefix = "ref_"

    def __init__(self, yaml_path, csv_path=None, arrays=None):
        if not have_ref():
            raise FileNotFoundError(REF_LIB + " missing: run `make -C oracle ref` where /root/reference exists")
        type(self).lib = type(self).lib or ct.CDLL(REF_LIB)
        self._bind()
        if arrays is None:
            f = self._f("create", ct.c_void_p, [ct.c_char_p, ct.c_char_p])
            self.h = f(yaml_path.encode(), (csv_path or "").encode())
        else:
            t, p, v = arrays
            t = np.ascontiguousarray(t, dtype=np.int32)
            p = np.ascontiguousarray(p, dtype=np.float64)
            v = np.ascontiguousarray(v, dtype=np.float64)
            f = self._f("create_from_arrays", ct.c_void_p,
                        [ct.c_char_p, ct.c_int64, ct.c_void_p, ct.c_void_p, ct.c_void_p])
            self.h = f(yaml_path.encode(), len(t), t.ctypes.data, p.ctypes.data, v.ctypes.data)


class OracleMPS(_Base):
    """Our CPU restatement (oracle/mps_oracle.cpp). Takes the flattened settings POD."""

    prefix = "orc_"

    def __init__(self, csettings: CSettings, types, pos, vel):
        if not have_oracle():
            raise FileNotFoundError(ORACLE_LIB + " missing: run `make -C oracle oracle`")
        type(self).lib = type(self).lib or ct.CDLL(ORACLE_LIB)
        self._bind()
        t = np.ascontiguousarray(types, dtype=np.int32)
        p = np.ascontiguousarray(pos, dtype=np.float64)
        v = np.ascontiguousarray(vel, dtype=np.float64)
        f = self._f("create", ct.c_void_p, [ct.POINTER(CSettings), ct.c_int64, ct.c_void_p, ct.c_void_p, ct.c_void_p])
        self.h = f(ct.byref(csettings), len(t), t.ctypes.data, p.ctypes.data, v.ctypes.data)
