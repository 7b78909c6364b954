"""The allocation-free entry point of the C ABI: mpsb_courant_static (csrc/k_static_scratch.cu).

Some users of the library cannot let it allocate device memory when they call it (all device memory already belongs
to their own pool; cudaMalloc's device-wide synchronisation would stall their other streams; a watchdog that only
holds device pointers to a live state). For them the library computes the Courant number of a particle state from
caller-owned device arrays with scratch memory that is a fixed array inside the library image instead of an
allocation. This file checks, through the C ABI, that the value is the one the handle-based entry point
(mpsb_run_phase with the Courant phase) gives for the same state — bit for bit, as the other parity tests ask of
every non-Krylov phase — on two calls in a row that share the scratch array: the second is smaller and slower than
the first, so anything the first left behind in the array would show.

The file name sorts last on purpose: pytest -x stops at the first failure, and a failure here must not hide the
parity tests. It does not depend on them: it passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as ct

import numpy as np
import pytest

from mps_b200 import capi, cases as C

pytestmark = pytest.mark.gpu

MAX_PARTICLES = 67108864  # MPSB_STATIC_MAX_PARTICLES in include/mps_b200.h


def _state(nx, seed, speed):
    """A scaled dam break with seeded velocities on EVERY particle (walls too, faster than the fluid, and a few
    ghosts faster still: only fluid particles may count)."""
    S, t, p, _ = C.scaled_dambreak(nx)
    rng = np.random.Generator(np.random.PCG64(seed))
    v = np.zeros_like(p)
    v[:, :2] = (rng.random((len(t), 2)) - 0.5) * speed
    v[t != C.FLUID] *= 3.0
    t = t.copy()
    fluid = np.flatnonzero(t == C.FLUID)
    t[fluid[:: max(1, len(fluid) // 7)]] = C.GHOST
    v[t == C.GHOST] *= 10.0
    return S, t, p, v


def _handle_courant(S, t, p, v):
    """The existing entry point: a handle, its cell-ordered copy of the state, the Courant phase."""
    dev = capi.DeviceMPS(S, t, p, v)
    dev.run_phase(capi.PH_NEIGHBORS)
    c = dev.run_phase(capi.PH_COURANT).courant
    dev.close()
    return c


def _static_courant(lib, torch, S, t, v, out):
    """The allocation-free entry point on device arrays this test owns (torch allocates them; the library does not)."""
    d_t = torch.from_numpy(np.ascontiguousarray(t, dtype=np.int32)).cuda()
    d_v = torch.from_numpy(np.ascontiguousarray(v, dtype=np.float64)).cuda()
    err = ct.create_string_buffer(256)
    rc = lib.mpsb_courant_static(d_t.data_ptr(), d_v.data_ptr(), len(t), S.dt, S.particle_distance, 0,
                                 None, out.data_ptr(), err, len(err))
    assert rc == 0, err.value.decode()
    torch.cuda.synchronize()  # the call only queues the kernel
    return float(out.item())


def test_courant_static_matches_the_handle_entry_point_on_two_calls_in_a_row():
    import torch

    lib = capi.load_library()
    out = torch.full((1,), -1.0, dtype=torch.float64, device="cuda")

    big = _state(400, seed=11, speed=0.8)       # 84 518 particles -> 42 thread blocks
    small = _state(64, seed=12, speed=0.05)     # fewer blocks, and every value below the first call's maximum
    want_big, want_small = _handle_courant(*big), _handle_courant(*small)
    assert want_big > want_small > 0.0

    S, t, _, v = big
    got_big = _static_courant(lib, torch, S, t, v, out)
    S, t, _, v = small
    got_small = _static_courant(lib, torch, S, t, v, out)
    assert got_big == want_big      # bit-exact: a maximum does not depend on the order it is taken in
    assert got_small == want_small  # nothing of the first call survived in the shared scratch array

    # and the plain definition, reference src/mps.cpp:547-571, on the host
    for (S, t, _, v), got in ((big, got_big), (small, got_small)):
        f = t == C.FLUID
        u = v[f]
        ref = np.max(np.sqrt((u[:, 0] * u[:, 0] + u[:, 1] * u[:, 1]) + u[:, 2] * u[:, 2]) * S.dt / S.particle_distance)
        assert got == float(ref)

    # no fluid at all, and no particles at all: 0.0, as the reference's loop leaves its initial value
    S, t, _, v = small
    assert _static_courant(lib, torch, S, np.full_like(t, C.WALL), v, out) == 0.0
    assert _static_courant(lib, torch, S, t[:0], v[:0], out) == 0.0


def test_courant_static_refuses_more_particles_than_its_scratch_serves():
    import torch

    lib = capi.load_library()
    out = torch.full((1,), -1.0, dtype=torch.float64, device="cuda")
    err = ct.create_string_buffer(256)
    # rejected on the host before any pointer is used: the inputs need not be that large
    rc = lib.mpsb_courant_static(out.data_ptr(), out.data_ptr(), MAX_PARTICLES + 1, 1e-3, 1e-2, 0, None,
                                 out.data_ptr(), err, len(err))
    assert rc == -1 and b"at most 67108864" in err.value
    torch.cuda.synchronize()
    assert float(out.item()) == -1.0  # nothing was launched
    assert lib.mpsb_courant_static(None, None, 5, 1e-3, 1e-2, 0, None, out.data_ptr(), None, 0) == -1
