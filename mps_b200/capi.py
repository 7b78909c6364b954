"""ctypes binding of libmps_b200.so — the C ABI declared in include/mps_b200.h.

`DeviceMPS` is the Python flavour of the host-side mirror of the reference's `MPS` class
(src/mps.hpp:22-71): `load_initial_state()` / `step_forward(is_time_to_export)` with the same
meaning, on top of the C ABI. There is no CPU path: if the library is missing or no CUDA
device is present the constructor raises.
"""
from __future__ import annotations

import ctypes as ct
import os

import numpy as np

from .cases import CSettings, Settings

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MPSB_LIB_PATH") or os.path.join(HERE, "libmps_b200.so")  # override: experiment builds

(PH_NEIGHBORS, PH_EXPLICIT, PH_COLLISION, PH_FLAGS_SOURCE, PH_MATRIX, PH_DIRICHLET, PH_SOLVE, PH_PRESSURE_POST,
 PH_GRADIENT_MOVE, PH_NUMBER_DENSITY, PH_COURANT) = range(11)

F_TYPE, F_POSITION, F_VELOCITY, F_ACCEL, F_PRESSURE, F_MIN_PRESSURE, F_RATIO, F_BC = (1 << k for k in range(8))
F_ALL = 0xFF

EXPORTED = [
    "mpsb_create", "mpsb_destroy", "mpsb_last_error", "mpsb_version", "mpsb_step", "mpsb_run_phase",
    "mpsb_upload", "mpsb_download", "mpsb_num_particles", "mpsb_get_neighbors", "mpsb_get_ppe", "mpsb_events",
    "mpsb_comm_unique_id", "mpsb_attach_comm", "mpsb_upload_device", "mpsb_launch_count", "mpsb_set_stream",
    "mpsb_debug_cg_cycles", "mpsb_debug_cg_block_cycles", "mpsb_debug_spmv", "mpsb_host_alloc", "mpsb_host_free",
    "mpsb_upload_shard", "mpsb_download_shard", "mpsb_shard_range", "mpsb_courant_static",
]


class CStepStats(ct.Structure):
    _fields_ = [
        ("courant", ct.c_double), ("cg_rel_residual", ct.c_double), ("ms_total", ct.c_double),
        ("ms_neighbors", ct.c_double), ("ms_explicit", ct.c_double), ("ms_collision", ct.c_double),
        ("ms_assemble", ct.c_double), ("ms_solve", ct.c_double), ("ms_correct", ct.c_double),
        ("cg_iterations", ct.c_int32), ("n_active", ct.c_int32), ("n_inner", ct.c_int32), ("n_surface", ct.c_int32),
        ("nnz", ct.c_int32), ("n_ghosted", ct.c_int32), ("n_unconnected", ct.c_int32), ("n_collisions", ct.c_int32),
        ("max_neighbors", ct.c_int32), ("courant_exceeds_cfl", ct.c_int32), ("ppe_columns_16bit", ct.c_int32), ("solve_in_cluster", ct.c_int32),
        ("ppe_matrix_free", ct.c_int32), ("cg_single_reduction", ct.c_int32),
    ]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class CFields(ct.Structure):
    _fields_ = [(n, ct.c_void_p) for n in ("type", "position", "velocity", "acceleration", "pressure",
                                            "minimum_pressure", "number_density_ratio", "boundary_condition")]


_FIELD_SPECS = [
    ("type", np.int32, 1, F_TYPE), ("position", np.float64, 3, F_POSITION), ("velocity", np.float64, 3, F_VELOCITY),
    ("acceleration", np.float64, 3, F_ACCEL), ("pressure", np.float64, 1, F_PRESSURE),
    ("minimum_pressure", np.float64, 1, F_MIN_PRESSURE), ("number_density_ratio", np.float64, 1, F_RATIO),
    ("boundary_condition", np.int32, 1, F_BC),
]

_lib = None


class MpsError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"mps_b200 error {code}: {msg}")
        self.code = code


def load_library():
    """Loads libmps_b200.so (must have been built in-tree: python -m mps_b200.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(f"{LIB_PATH} not built; run `python -m mps_b200.build` (there is no CPU fallback)")
    lib = ct.CDLL(LIB_PATH)
    vp, i64, i32 = ct.c_void_p, ct.c_int64, ct.c_int
    lib.mpsb_create.restype = i32
    lib.mpsb_create.argtypes = [ct.POINTER(CSettings), i64, vp, vp, vp, i32, ct.POINTER(vp)]
    lib.mpsb_destroy.restype = i32
    lib.mpsb_destroy.argtypes = [vp]
    lib.mpsb_last_error.restype = ct.c_char_p
    lib.mpsb_last_error.argtypes = [vp]
    lib.mpsb_version.restype = ct.c_char_p
    lib.mpsb_step.restype = i32
    lib.mpsb_step.argtypes = [vp, i32, ct.POINTER(CStepStats)]
    lib.mpsb_run_phase.restype = i32
    lib.mpsb_run_phase.argtypes = [vp, i32, ct.POINTER(CStepStats)]
    lib.mpsb_upload.restype = i32
    lib.mpsb_upload.argtypes = [vp, ct.c_uint32, ct.POINTER(CFields)]
    lib.mpsb_download.restype = i32
    lib.mpsb_download.argtypes = [vp, ct.c_uint32, ct.POINTER(CFields)]
    lib.mpsb_num_particles.restype = i64
    lib.mpsb_num_particles.argtypes = [vp]
    lib.mpsb_get_neighbors.restype = i32
    lib.mpsb_get_neighbors.argtypes = [vp, vp, vp, vp, i64, ct.POINTER(i64)]
    lib.mpsb_get_ppe.restype = i32
    lib.mpsb_get_ppe.argtypes = [vp, vp, vp, vp, vp, i64, ct.POINTER(i64)]
    lib.mpsb_events.restype = i32
    lib.mpsb_events.argtypes = [vp, i32, vp, vp, i32, ct.POINTER(i32)]
    lib.mpsb_comm_unique_id.restype = i32
    lib.mpsb_comm_unique_id.argtypes = [vp]
    lib.mpsb_attach_comm.restype = i32
    lib.mpsb_attach_comm.argtypes = [vp, i32, i32, vp]
    for fn in (lib.mpsb_upload_shard, lib.mpsb_download_shard):
        fn.restype = i32
        fn.argtypes = [vp, ct.c_uint32, ct.POINTER(CFields)]
    lib.mpsb_shard_range.restype = i32
    lib.mpsb_shard_range.argtypes = [vp, ct.POINTER(ct.c_int64), ct.POINTER(ct.c_int64)]
    lib.mpsb_upload_device.restype = i32
    lib.mpsb_upload_device.argtypes = [vp, vp, vp, vp]
    lib.mpsb_set_stream.restype = i32
    lib.mpsb_set_stream.argtypes = [vp, vp]
    lib.mpsb_launch_count.restype = i64
    lib.mpsb_launch_count.argtypes = [vp]
    lib.mpsb_courant_static.restype = i32
    lib.mpsb_courant_static.argtypes = [vp, vp, i64, ct.c_double, ct.c_double, i32, vp, vp, ct.c_char_p, i32]
    _lib = lib
    return lib


def new_unique_id(lib=None):
    """128-byte ncclUniqueId from mpsb_comm_unique_id (rank 0 only)."""
    lib = lib or load_library()
    buf = (ct.c_char * 128)()
    rc = lib.mpsb_comm_unique_id(buf)
    if rc != 0:
        raise MpsError(rc, "mpsb_comm_unique_id failed (libnccl.so.2 not loadable?)")
    return bytes(buf.raw)


def distribute_unique_id(rank, make_id, group=None):
    """The id made by `make_id()` on rank 0, on every rank of the torch.distributed group."""
    import torch
    import torch.distributed as dist

    box = [make_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group, device=torch.device("cpu") if dist.get_backend(group) == "gloo" else None)
    uid = box[0]
    if not isinstance(uid, (bytes, bytearray)) or len(uid) != 128:
        raise ValueError("unique id must be 128 bytes")
    return bytes(uid)


def row_block(np_rows, rank, world):
    """Rows [lo, hi) of the padded sorted order that `rank` solves (k_pcg.cu: rows_per_rank = Np / world);
    np_rows = N rounded up to a multiple of 256 (mpsb_api.cu), so 1/2/4/8 ranks get whole 32-row slices."""
    if world not in (1, 2, 4, 8) or not 0 <= rank < world:
        raise ValueError("supported world sizes: 1, 2, 4, 8")
    if np_rows % 256:
        raise ValueError("np_rows must be a multiple of 256")
    per = np_rows // world
    return rank * per, (rank + 1) * per


def padded_rows(n):
    return (int(n) + 255) // 256 * 256


def shard_block(n, rank, world):
    """Particle ids [first, last) that the sharded transfers of `rank` move (mpsb_api.cu shard_block: the row-block split
    applied to ids and clipped to n); the blocks of all ranks partition [0, n)."""
    lo, hi = row_block(padded_rows(n), rank, world)
    return min(n, lo), min(n, hi)


class DeviceMPS:
    """One MPS simulation state resident on one B200."""

    def __init__(self, settings, types, pos, vel, device=0):
        self.lib = load_library()
        self.settings = settings
        self.cs = settings.to_c() if isinstance(settings, Settings) else settings
        t = np.ascontiguousarray(types, dtype=np.int32)
        p = np.ascontiguousarray(pos, dtype=np.float64)
        v = np.ascontiguousarray(vel, dtype=np.float64)
        self.n = len(t)
        self.h = ct.c_void_p()
        rc = self.lib.mpsb_create(ct.byref(self.cs), self.n, t.ctypes.data, p.ctypes.data, v.ctypes.data, device,
                                  ct.byref(self.h))
        if rc != 0:
            raise MpsError(rc, self.lib.mpsb_last_error(None).decode())
        self.debug_log_count = 0
        self.last_stats = None

    def close(self):
        if getattr(self, "h", None):
            self.lib.mpsb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise MpsError(rc, self.lib.mpsb_last_error(self.h).decode())

    # --- reference MPS surface -------------------------------------------------------------
    def load_initial_state(self):
        """The tail of MPS::loadInitialState (src/mps.cpp:106): setNumberDensity()."""
        return self.run_phase(PH_NUMBER_DENSITY)

    def step_forward(self, is_time_to_export=False):
        """MPS::stepForward (src/mps.cpp:111-130); returns the Courant number."""
        st = CStepStats()
        want = bool(is_time_to_export) or bool(self.cs.higher_order_source_on)
        self._check(self.lib.mpsb_step(self.h, int(want), ct.byref(st)))
        self.last_stats = st
        self.debug_log_count += st.n_ghosted + st.n_unconnected + st.courant_exceeds_cfl
        return st.courant

    # --- pieces ----------------------------------------------------------------------------------
    def run_phase(self, phase):
        st = CStepStats()
        self._check(self.lib.mpsb_run_phase(self.h, int(phase), ct.byref(st)))
        self.last_stats = st
        return st

    def download(self, names=None):
        arrs, cf, mask = {}, CFields(), 0
        for name, dt, w, bit in _FIELD_SPECS:
            if names is not None and name not in names:
                continue
            a = np.zeros((self.n, w) if w > 1 else self.n, dtype=dt)
            arrs[name] = a
            setattr(cf, name, a.ctypes.data)
            mask |= bit
        self._check(self.lib.mpsb_download(self.h, mask, ct.byref(cf)))
        return arrs

    def shard_range(self):
        """[first, last): the block of particle ids the sharded transfers of this rank move."""
        lo, hi = ct.c_int64(), ct.c_int64()
        self._check(self.lib.mpsb_shard_range(self.h, ct.byref(lo), ct.byref(hi)))
        return int(lo.value), int(hi.value)

    def download_into(self, arrs, shard=False):
        """Download into caller-owned (e.g. pinned) contiguous numpy arrays, keyed by field name.
        shard=True (several ranks, collective): only this rank's block of ids is written."""
        cf, mask = CFields(), 0
        for name, dt, w, bit in _FIELD_SPECS:
            a = arrs.get(name)
            if a is None:
                continue
            if a.dtype != dt or not a.flags.c_contiguous or a.size != self.n * w:
                raise ValueError(f"download_into: {name} must be C-contiguous {dt.__name__}[{self.n}*{w}]")
            setattr(cf, name, a.ctypes.data)
            mask |= bit
        self._check((self.lib.mpsb_download_shard if shard else self.lib.mpsb_download)(self.h, mask, ct.byref(cf)))
        return arrs

    def upload(self, arrs, shard=False):
        """shard=True (several ranks, collective): only this rank's block of ids crosses PCIe, the GPUs all-gather the rest."""
        cf, mask, keep = CFields(), 0, []
        for name, dt, w, bit in _FIELD_SPECS:
            if name in arrs and arrs[name] is not None:
                a = np.ascontiguousarray(arrs[name], dtype=dt)
                keep.append(a)
                setattr(cf, name, a.ctypes.data)
                mask |= bit
        self._check((self.lib.mpsb_upload_shard if shard else self.lib.mpsb_upload)(self.h, mask, ct.byref(cf)))

    def neighbors(self):
        start = np.zeros(self.n + 1, dtype=np.int64)
        total = ct.c_int64()
        self._check(self.lib.mpsb_get_neighbors(self.h, start.ctypes.data, None, None, 0, ct.byref(total)))
        ids = np.zeros(total.value, dtype=np.int32)
        dist = np.zeros(total.value, dtype=np.float64)
        self._check(self.lib.mpsb_get_neighbors(self.h, start.ctypes.data, ids.ctypes.data, dist.ctypes.data,
                                                total.value, ct.byref(total)))
        return start, ids, dist

    def ppe(self):
        start = np.zeros(self.n + 1, dtype=np.int64)
        rhs = np.zeros(self.n, dtype=np.float64)
        nnz = ct.c_int64()
        self._check(self.lib.mpsb_get_ppe(self.h, start.ctypes.data, None, None, rhs.ctypes.data, 0, ct.byref(nnz)))
        col = np.zeros(nnz.value, dtype=np.int32)
        val = np.zeros(nnz.value, dtype=np.float64)
        self._check(self.lib.mpsb_get_ppe(self.h, start.ctypes.data, col.ctypes.data, val.ctypes.data,
                                          rhs.ctypes.data, nnz.value, ct.byref(nnz)))
        return start, col, val, rhs

    def events(self, kind, capacity=4096):
        ids = np.zeros(capacity, dtype=np.int32)
        xyz = np.zeros((capacity, 3), dtype=np.float64)
        cnt = ct.c_int()
        self._check(self.lib.mpsb_events(self.h, kind, ids.ctypes.data, xyz.ctypes.data, capacity, ct.byref(cnt)))
        m = min(cnt.value, capacity)
        return ids[:m], xyz[:m]

    # --- multi-GPU (one process per GPU) ---------------------------------------------------------
    def attach_comm(self, rank, world, unique_id128):
        """Join the `world`-rank exchange group (include/mps_b200.h: mpsb_attach_comm)."""
        buf = (ct.c_char * 128).from_buffer_copy(bytes(unique_id128))
        self._check(self.lib.mpsb_attach_comm(self.h, int(rank), int(world), buf))
        self.rank, self.world = int(rank), int(world)

    def attach_comm_torch(self, rank, world, group=None):
        """Rank 0 makes the NCCL unique id, torch.distributed (any backend) hands it to the others."""
        uid = distribute_unique_id(rank, lambda: new_unique_id(self.lib), group)
        self.attach_comm(rank, world, uid)

    def set_stream(self, cuda_stream_ptr):
        """Run on the caller's CUDA stream (e.g. torch.cuda.current_stream().cuda_stream)."""
        self._check(self.lib.mpsb_set_stream(self.h, ct.c_void_p(cuda_stream_ptr)))

    def cg_cycles(self):
        out = (ct.c_int64 * 4)()
        self.lib.mpsb_debug_cg_cycles.argtypes = [ct.c_void_p, ct.c_void_p]
        self._check(self.lib.mpsb_debug_cg_cycles(self.h, out))
        return list(out)

    def cg_block_cycles(self):
        """(5, nblocks) int64: per-CTA cycles in sweep A, B, C, reductions+barriers, and the SM id."""
        cap = 1024
        out = np.zeros((5, cap), dtype=np.int64)
        nb = ct.c_int()
        self.lib.mpsb_debug_cg_block_cycles.argtypes = [ct.c_void_p, ct.c_void_p, ct.c_int, ct.POINTER(ct.c_int)]
        flat = np.zeros(5 * cap, dtype=np.int64)
        self._check(self.lib.mpsb_debug_cg_block_cycles(self.h, flat.ctypes.data, cap, ct.byref(nb)))
        return flat[:5 * nb.value].reshape(5, nb.value)

    def spmv(self, x, matrix_free):
        """y = A x with the last assembled PPE (vectors by particle id); stored entries or the matrix-free operator."""
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.zeros_like(x)
        self.lib.mpsb_debug_spmv.restype = ct.c_int
        self.lib.mpsb_debug_spmv.argtypes = [ct.c_void_p, ct.c_void_p, ct.c_void_p, ct.c_int]
        self._check(self.lib.mpsb_debug_spmv(self.h, x.ctypes.data, y.ctypes.data, int(matrix_free)))
        return y

    @property
    def launch_count(self):
        return int(self.lib.mpsb_launch_count(self.h))
