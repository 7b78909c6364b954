"""Case inputs for the MPS step: settings (YAML), initial lattices (CSV or in-memory), RefValues.

Host-side mirror (Python flavour, for tests and bench.py; the C++ flavour lives in
mps_b200/host/) of
  * reference src/settings.cpp:12-89   (YAML keys -> Settings)
  * reference src/mps.cpp:132-163      (RefValues: n0, lambda with the same loop order)
  * reference inputs/dambreak/main.cpp:26-113 (lattice generator, x-outer / y-inner ids)
  * reference src/exporter.cpp:10-56   (input.csv layout) and src/mps.cpp:38-94 (its reader)
"""
from __future__ import annotations

import ctypes
import math
import os
from dataclasses import dataclass, field

import numpy as np

REPO_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES_DIR = os.path.join(REPO_ROOT, "cases")

GHOST, FLUID, WALL, DUMMY_WALL = -1, 0, 1, 2
BC_IGNORED, BC_SURFACE, BC_INNER = 0, 1, 2


class CSettings(ctypes.Structure):
    """ctypes image of `mpsb_settings` (include/mps_b200.h)."""

    _fields_ = [
        ("dim", ctypes.c_int32),
        ("quasi_compressibility_on", ctypes.c_int32),
        ("higher_order_source_on", ctypes.c_int32),
        ("reserved0", ctypes.c_int32),
        ("particle_distance", ctypes.c_double),
        ("dt", ctypes.c_double),
        ("cfl_condition", ctypes.c_double),
        ("domain_min", ctypes.c_double * 3),
        ("domain_max", ctypes.c_double * 3),
        ("density", ctypes.c_double),
        ("kinematic_viscosity", ctypes.c_double),
        ("gravity", ctypes.c_double * 3),
        ("re_pressure", ctypes.c_double),
        ("re_viscosity", ctypes.c_double),
        ("re_surface", ctypes.c_double),
        ("re_max", ctypes.c_double),
        ("surface_threshold", ctypes.c_double),
        ("compressibility", ctypes.c_double),
        ("higher_order_gamma", ctypes.c_double),
        ("relaxation_coefficient", ctypes.c_double),
        ("collision_distance", ctypes.c_double),
        ("restitution", ctypes.c_double),
        ("n0_pressure", ctypes.c_double),
        ("lambda_pressure", ctypes.c_double),
        ("n0_viscosity", ctypes.c_double),
        ("lambda_viscosity", ctypes.c_double),
        ("n0_surface", ctypes.c_double),
        ("lambda_surface", ctypes.c_double),
    ]


def ref_values(dim: int, l0: float, re: float) -> tuple[float, float]:
    """(n0, lambda) exactly as reference RefValues::RefValues (src/mps.cpp:132-163):
    iX outer, iY, iZ inner; dist = sqrt((x*x+y*y)+z*z); dist2 = dist*dist; lambda /= n0."""
    n0 = 0.0
    lam = 0.0
    zr = range(0, 1) if dim == 2 else range(-4, 5)
    for ix in range(-4, 5):
        for iy in range(-4, 5):
            for iz in zr:
                if ix == 0 and iy == 0 and iz == 0:
                    continue
                x, y, z = l0 * float(ix), l0 * float(iy), l0 * float(iz)
                dist = math.sqrt((x * x + y * y) + z * z)
                dist2 = dist * dist
                w = (re / dist) - 1.0 if dist < re else 0.0
                n0 += w
                lam += dist2 * w
    lam /= n0
    return n0, lam


@dataclass
class Settings:
    """Field-for-field reference `Settings` (src/settings.hpp:35-86); radii already times l0."""

    dim: int = 2
    particle_distance: float = 0.0
    dt: float = 0.0
    finish_time: float = 0.0
    output_interval: float = 0.0
    cfl_condition: float = 0.0
    domain_min: tuple = (0.0, 0.0, 0.0)
    domain_max: tuple = (0.0, 0.0, 0.0)
    density: float = 0.0
    kinematic_viscosity: float = 0.0
    gravity: tuple = (0.0, 0.0, 0.0)
    re_pressure: float = 0.0
    re_viscosity: float = 0.0
    re_surface: float = 0.0
    re_max: float = 0.0
    surface_threshold: float = 0.0
    quasi_compressibility_on: bool = False
    compressibility: float = 0.0
    higher_order_source_on: bool = False
    higher_order_gamma: float = 0.0
    relaxation_coefficient: float = 0.0
    collision_distance: float = 0.0
    restitution: float = 0.0
    input_csv_path: str = ""
    extras: dict = field(default_factory=dict)

    def to_c(self) -> CSettings:
        c = CSettings()
        c.dim = self.dim
        c.quasi_compressibility_on = int(self.quasi_compressibility_on)
        c.higher_order_source_on = int(self.higher_order_source_on)
        c.particle_distance = self.particle_distance
        c.dt = self.dt
        c.cfl_condition = self.cfl_condition
        for k in range(3):
            c.domain_min[k] = self.domain_min[k]
            c.domain_max[k] = self.domain_max[k]
            c.gravity[k] = self.gravity[k]
        c.density = self.density
        c.kinematic_viscosity = self.kinematic_viscosity
        c.re_pressure = self.re_pressure
        c.re_viscosity = self.re_viscosity
        c.re_surface = self.re_surface
        c.re_max = self.re_max
        c.surface_threshold = self.surface_threshold
        c.compressibility = self.compressibility if self.quasi_compressibility_on else 0.0
        c.higher_order_gamma = self.higher_order_gamma if self.higher_order_source_on else 0.0
        c.relaxation_coefficient = self.relaxation_coefficient
        c.collision_distance = self.collision_distance
        c.restitution = self.restitution
        c.n0_pressure, c.lambda_pressure = ref_values(self.dim, self.particle_distance, self.re_pressure)
        c.n0_viscosity, c.lambda_viscosity = ref_values(self.dim, self.particle_distance, self.re_viscosity)
        c.n0_surface, c.lambda_surface = ref_values(self.dim, self.particle_distance, self.re_surface)
        return c


def load_settings(yaml_path: str) -> Settings:
    """Same keys and derived values as reference Settings::load (src/settings.cpp:20-84)."""
    import yaml

    with open(yaml_path) as f:
        root = yaml.safe_load(f)
    s = Settings()
    s.dim = int(root["dim"])
    l0 = s.particle_distance = float(root["particle distance"])
    s.dt = float(root["dt"])
    s.finish_time = float(root["finish time"])
    s.output_interval = float(root["output interval"])
    s.cfl_condition = float(root["CFL condition"])
    dx, dy, dz = root["domain range X"], root["domain range Y"], root["domain range Z"]
    s.domain_min = (float(dx[0]), float(dy[0]), float(dz[0]))
    s.domain_max = (float(dx[1]), float(dy[1]), float(dz[1]))
    s.density = float(root["density"])
    s.kinematic_viscosity = float(root["kinematic viscosity"])
    s.gravity = tuple(float(g) for g in root["gravity"])
    rr = root["effective radius ratio"]
    s.re_pressure = float(rr["pressure"]) * l0
    s.re_viscosity = float(rr["viscosity"]) * l0
    s.re_surface = float(rr["surface detection"]) * l0
    s.re_max = max(s.re_pressure, s.re_viscosity, s.re_surface)
    s.surface_threshold = float(root["threshold for surface detection"])
    ps = root["pressure stabilization"]

    def _on(node):  # PyYAML (YAML 1.1) reads the bare key `on` as boolean True
        return bool(node["on"] if "on" in node else node[True])

    s.quasi_compressibility_on = _on(ps["quasi compressibility"])
    if s.quasi_compressibility_on:
        s.compressibility = float(ps["quasi compressibility"]["compressibility"])
    s.higher_order_source_on = _on(ps["higher order source term"])
    if s.higher_order_source_on:
        s.higher_order_gamma = float(ps["higher order source term"]["gamma"])
    s.relaxation_coefficient = float(root["relaxation coefficient for pressure"])
    s.collision_distance = float(root["collision distance ratio"]) * l0
    s.restitution = float(root["coefficient of restitution"])
    s.input_csv_path = str(root["inputCsvPath"])
    return s


# Geometry constants of the three shipped generators (reference inputs/<case>/main.cpp:12-19).
GEOMETRIES = {
    "dambreak": dict(l0=0.01, domain_x=(0.0, 0.3), domain_y=(0.0, 0.25), fluid_x=(0.0, 0.1), fluid_y=(0.0, 0.2)),
    "dambreak_fine": dict(l0=0.004, domain_x=(0.0, 1.0), domain_y=(0.0, 0.25), fluid_x=(0.0, 0.4), fluid_y=(0.0, 0.2)),
    "hydrostatic": dict(l0=0.005, domain_x=(0.0, 0.2), domain_y=(0.0, 0.35), fluid_x=(0.0, 0.2), fluid_y=(0.0, 0.3)),
}


def _c_round(x: float) -> int:
    """C `round` (half away from zero), as used by the generator (inputs/dambreak/main.cpp:33-36)."""
    return int(math.floor(abs(x) + 0.5)) * (1 if x >= 0 else -1)


def generate_lattice(l0, domain_x, domain_y, fluid_x, fluid_y, wall_layer=1, dummy_wall_layer=2):
    """2-D lattice of the reference generators (inputs/dambreak/main.cpp:26-80): fluid box, wall
    layer, dummy-wall layers, open top; ids dense in loop order x outer / y inner. Returns
    (type int32[n], position float64[n,3]) with position = ((ix+0.5)*l0, (iy+0.5)*l0, 0) in full
    double precision (the shipped CSVs hold these rounded to 6 decimals)."""
    nx = (fluid_x[1] - fluid_x[0]) / l0
    ny = (fluid_y[1] - fluid_y[0]) / l0
    eps = 0.01 * l0
    if abs(nx - round(nx)) > eps or abs(ny - round(ny)) > eps:
        raise ValueError("fluid range is not divisible by the particle distance")  # main.cpp:82-101
    a = dummy_wall_layer + wall_layer
    ix = np.arange(_c_round(domain_x[0] / l0) - a, _c_round(domain_x[1] / l0) + a + 1)
    iy = np.arange(_c_round(domain_y[0] / l0) - a, _c_round(domain_y[1] / l0) + a + 1)
    x = ((ix + 0.5) * l0)[:, None] + np.zeros(len(iy))[None, :]
    y = np.zeros(len(ix))[:, None] + ((iy + 0.5) * l0)[None, :]
    x_max, y_max = domain_x[1], domain_y[1]

    def inside(xr, yr):
        xr, yr = sorted(xr), sorted(yr)
        return (xr[0] - eps < x) & (x < xr[1] + eps) & (yr[0] - eps < y) & (y < yr[1] + eps)

    t = np.full(x.shape, GHOST, dtype=np.int32)
    t[inside((-a * l0, x_max + a * l0), (-a * l0, y_max))] = DUMMY_WALL
    t[inside((-wall_layer * l0, x_max + wall_layer * l0), (-wall_layer * l0, y_max))] = WALL
    t[inside((0.0, x_max), (0.0, y_max))] = GHOST
    t[inside(tuple(fluid_x), tuple(fluid_y))] = FLUID
    keep = (t != GHOST).ravel()  # row-major ravel == x outer, y inner
    pos = np.zeros((int(keep.sum()), 3))
    pos[:, 0] = x.ravel()[keep]
    pos[:, 1] = y.ravel()[keep]
    return np.ascontiguousarray(t.ravel()[keep]), pos


def write_input_csv(path, types, pos, vel=None, time=0.0, pressure=None, ratio=None):
    """input.csv in the layout of reference Exporter::toCsv (src/exporter.cpp:10-56) with the
    `%f` float form of the shipped files (clang build of csv-parser). Ghosts skipped."""
    n = len(types)
    vel = np.zeros((n, 3)) if vel is None else vel
    pressure = np.zeros(n) if pressure is None else pressure
    ratio = np.zeros(n) if ratio is None else ratio
    with open(path, "w", newline="") as f:
        f.write("Time (s),%f\n" % time)
        f.write("Number of Particles,%d\n" % n)
        f.write(
            "ID,Type,Position.x (m),Position.y (m),Position.z (m),Velocity.x (m/s),Velocity.y (m/s),"
            "Velocity.z (m/s),Pressure (Pa),Number Density Ratio\n"
        )
        for i in range(n):
            if types[i] == GHOST:
                continue
            f.write(
                "%d,%d,%f,%f,%f,%f,%f,%f,%f,%f\n"
                % (i, types[i], pos[i, 0], pos[i, 1], pos[i, 2], vel[i, 0], vel[i, 1], vel[i, 2], pressure[i], ratio[i])
            )


def load_input_csv(path):
    """Reader with the semantics of reference MPS::importInitialCondition (src/mps.cpp:38-94):
    time from row 1 col 2, header on row 3, columns ID/Type/Position/Velocity; pressure and
    number-density columns ignored. Returns (time, type int32[n], pos[n,3], vel[n,3])."""
    with open(path) as f:
        lines = f.read().splitlines()
    time = float(lines[0].split(",")[1])
    header = lines[2].split(",")
    col = {name: k for k, name in enumerate(header)}
    rows = [ln.split(",") for ln in lines[3:] if ln.strip()]
    n = len(rows)
    types = np.empty(n, dtype=np.int32)
    pos = np.empty((n, 3))
    vel = np.empty((n, 3))
    for i, r in enumerate(rows):
        types[i] = int(r[col["Type"]])
        pos[i] = (float(r[col["Position.x (m)"]]), float(r[col["Position.y (m)"]]), float(r[col["Position.z (m)"]]))
        vel[i] = (float(r[col["Velocity.x (m/s)"]]), float(r[col["Velocity.y (m/s)"]]), float(r[col["Velocity.z (m/s)"]]))
    return time, types, pos, vel


def case_paths(name: str):
    d = os.path.join(CASES_DIR, name)
    return os.path.join(d, "settings.yml"), os.path.join(d, "input.csv")


def load_case(name: str):
    """(Settings, type, pos, vel) of a shipped case under cases/<name>/ (or the generated 3-D cases "tank3d" / "tank3d_r31")."""
    if name == "tank3d":
        return tank3d(jitter=0.05)
    if name == "tank3d_r31":  # the radius ratio every settings.yml of the reference ships (tutorials/dambreak/settings.yml:27-30): ~120 neighbours in 3-D
        return tank3d(ratio=3.1, jitter=0.05)
    yml, _ = case_paths(name)
    s = load_settings(yml)
    csv = s.input_csv_path if os.path.isabs(s.input_csv_path) else os.path.join(REPO_ROOT, s.input_csv_path)
    _, t, p, v = load_input_csv(csv)
    return s, t, p, v


def scaled_dambreak(nx_fluid: int, base: str = "dambreak_fine"):
    """Synthetic scaled dam break (BASELINE.json configs[3]): dambreak_fine
    geometry with l0 = 0.4/nx_fluid (nx_fluid a multiple of 8 keeps all box edges integral), dt
    scaled to keep dt/l0 = 0.125 s/m, every other setting as cases/dambreak_fine/settings.yml.
    Generated in memory in full double precision. nx_fluid = 2832 -> 4 010 112 fluid particles."""
    s = load_settings(case_paths(base)[0])
    g = GEOMETRIES[base]
    l0 = (g["fluid_x"][1] - g["fluid_x"][0]) / nx_fluid
    ratio_p, ratio_v, ratio_s = (s.re_pressure / s.particle_distance, s.re_viscosity / s.particle_distance,
                                 s.re_surface / s.particle_distance)
    col_ratio = s.collision_distance / s.particle_distance
    dt_ratio = s.dt / s.particle_distance
    s.particle_distance = l0
    s.dt = dt_ratio * l0
    s.re_pressure, s.re_viscosity, s.re_surface = ratio_p * l0, ratio_v * l0, ratio_s * l0
    s.re_max = max(s.re_pressure, s.re_viscosity, s.re_surface)
    s.collision_distance = col_ratio * l0
    s.input_csv_path = ""
    t, p = generate_lattice(l0, g["domain_x"], g["domain_y"], g["fluid_x"], g["fluid_y"])
    return s, t, p, np.zeros_like(p)


def jittered(s: Settings, types, pos, seed: int, amplitude: float = 0.1):
    """A mid-flow-like 2-D state for parity fixtures that cannot carry a state (tests/golden/large_*):
    fluid positions moved by a seeded uniform jitter of +-amplitude*l0 in x and y, fluid velocities from a
    polynomial field. Only +, -, * on doubles and numpy's PCG64 stream, so every machine regenerates the
    same bits from (nx, seed). Minimum distance stays (1 - 2*amplitude)*l0, above the collision distance."""
    rng = np.random.Generator(np.random.PCG64(seed))
    fluid = types == FLUID
    l0 = s.particle_distance
    p = pos.copy()
    d = (rng.random((len(types), 2)) - 0.5) * (2.0 * amplitude * l0)
    p[fluid, 0] += d[fluid, 0]
    p[fluid, 1] += d[fluid, 1]
    v = np.zeros_like(p)
    x, y = p[:, 0], p[:, 1]
    v[fluid, 0] = (0.3 * (x * (1.0 - x)) * (y - 0.3))[fluid]
    v[fluid, 1] = (-0.2 * (x * y))[fluid]
    return p, v


def write_settings_yaml(s: Settings, path: str, finish_time=None, output_interval=None):
    """settings.yml with the reference's exact keys (src/settings.cpp:20-84) for a generated case, so the
    compiled reference (whose MPS constructor only takes a YAML path) can run the same settings.
    Ratios are written with repr() so re = ratio*l0 round-trips bit-for-bit where the ratio is exact."""
    l0 = s.particle_distance

    def ratio(x):
        r = x / l0
        for digits in range(1, 17):  # shortest decimal that reproduces x when multiplied back
            c = float(f"{r:.{digits}g}")
            if c * l0 == x:
                return repr(_F(c))
        return repr(_F(r))

    yn = lambda b: "True" if b else "False"  # noqa: E731

    class _F(float):  # fkYAML reads `1e-06` as a non-float: always write a decimal point in the mantissa
        def __repr__(self):
            r = float.__repr__(self)
            if "e" in r and "." not in r:
                m, e = r.split("e")
                r = m + ".0e" + e
            return r
    lines = [
        "dim: %d" % s.dim,
        "particle distance: %r" % _F(l0),
        "dt: %r" % _F(s.dt),
        "finish time: %r" % _F(s.finish_time if finish_time is None else finish_time),
        "output interval: %r" % _F(s.output_interval if output_interval is None else output_interval),
        "CFL condition: %r" % _F(s.cfl_condition),
        "domain range X: [%r, %r]" % (_F(s.domain_min[0]), _F(s.domain_max[0])),
        "domain range Y: [%r, %r]" % (_F(s.domain_min[1]), _F(s.domain_max[1])),
        "domain range Z: [%r, %r]" % (_F(s.domain_min[2]), _F(s.domain_max[2])),
        "density: %r" % _F(s.density),
        "kinematic viscosity: %r" % _F(s.kinematic_viscosity),
        "gravity: [%r, %r, %r]" % tuple(_F(g) for g in s.gravity),
        "effective radius ratio:",
        "  pressure: %s" % ratio(s.re_pressure),
        "  viscosity: %s" % ratio(s.re_viscosity),
        "  surface detection: %s" % ratio(s.re_surface),
        "threshold for surface detection: %r" % _F(s.surface_threshold),
        "pressure stabilization:",
        "  quasi compressibility:",
        "    on: %s" % yn(s.quasi_compressibility_on),
        "    compressibility: %r" % _F(s.compressibility if s.quasi_compressibility_on else 0.0),
        "  higher order source term:",
        "    on: %s" % yn(s.higher_order_source_on),
        "    gamma: %r" % _F(s.higher_order_gamma if s.higher_order_source_on else 0.0),
        "relaxation coefficient for pressure: %r" % _F(s.relaxation_coefficient),
        "collision distance ratio: %s" % ratio(s.collision_distance),
        "coefficient of restitution: %r" % _F(s.restitution),
        "inputCsvPath: %s" % (s.input_csv_path or "none.csv"),
    ]
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return path


def tank3d(n_fluid=(8, 10, 6), n_tank=(16, 14, 6), l0=0.01, ratio=2.1, jitter=0.0):
    """Synthetic 3-D case (the reference ships none): an open-top tank of n_tank lattice cells
    with one wall layer and two dummy-wall layers on the bottom and the four sides, a fluid column of
    n_fluid cells in the corner x = z = 0, dambreak's physical settings with dim = 3 and radius ratio `ratio`.
    ids dense in loop order x outer, y, z inner (the reference generators' convention extended by one axis).
    `jitter` > 0 displaces fluid particles by a fixed pseudo-random pattern (mt19937 seed 1, +-jitter*l0) so
    that neighbour distances are not lattice-degenerate."""
    s = load_settings(case_paths("dambreak")[0])
    s.dim = 3
    s.particle_distance = l0
    s.dt = 0.1 * l0
    s.re_pressure = s.re_viscosity = s.re_surface = s.re_max = ratio * l0
    s.collision_distance = 0.5 * l0
    X, Y, Z = (n * l0 for n in n_tank)
    s.domain_min = (-0.1, -0.1, -0.1)
    s.domain_max = (X + 0.1, Y + 0.1, Z + 0.1)
    s.input_csv_path = ""
    a = 3
    ix = np.arange(-a, n_tank[0] + a)
    iy = np.arange(-a, n_tank[1])
    iz = np.arange(-a, n_tank[2] + a)
    gx, gy, gz = np.meshgrid(ix, iy, iz, indexing="ij")
    inside = lambda lo: (gx >= -lo) & (gx < n_tank[0] + lo) & (gy >= -lo) & (gz >= -lo) & (gz < n_tank[2] + lo)  # noqa: E731
    t = np.full(gx.shape, GHOST, dtype=np.int32)
    t[inside(3)] = DUMMY_WALL
    t[inside(1)] = WALL
    t[inside(0)] = GHOST
    t[(gx >= 0) & (gx < n_fluid[0]) & (gy >= 0) & (gy < n_fluid[1]) & (gz >= 0) & (gz < n_fluid[2])] = FLUID
    keep = (t != GHOST).ravel()
    pos = np.stack([(g.ravel()[keep] + 0.5) * l0 for g in (gx, gy, gz)], axis=1)
    types = np.ascontiguousarray(t.ravel()[keep])
    if jitter > 0.0:
        rng = np.random.Generator(np.random.MT19937(1))
        d = rng.uniform(-jitter * l0, jitter * l0, size=pos.shape)
        pos[types == FLUID] += d[types == FLUID]
    return s, types, np.ascontiguousarray(pos), np.zeros_like(pos)
