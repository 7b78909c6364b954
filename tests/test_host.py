"""Host C++ mirror (mps_b200/host): Settings/RefValues/CSV/VTU on CPU, and on the GPU box the whole CLI
`mps -s settings.yml` against the compiled reference binary (oracle/_ref/main.exe) file by file."""
import ctypes
import os
import shutil
import subprocess

import pytest

from mps_b200 import cases as C

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "mps_b200", "host")
SELFTEST = os.path.join(HOST, "host_selftest")
MPS_EXE = os.path.join(HOST, "mps")
GEN_EXE = os.path.join(HOST, "mps_generate")
REF_EXE = os.path.join(ROOT, "oracle", "_ref", "main.exe")
CASES = ["dambreak", "dambreak_fine", "hydrostatic", "dambreak_fine_higher_order", "dambreak_fine_original"]


@pytest.fixture(scope="module", autouse=True)
def _built():
    if not (os.path.exists(SELFTEST) and os.path.exists(MPS_EXE) and os.path.exists(GEN_EXE)):
        from mps_b200 import build as B

        B.build_library()
        B.build_host()
    assert os.path.exists(SELFTEST) and os.path.exists(MPS_EXE)


def _run(*args, **kw):
    return subprocess.run(list(args), cwd=ROOT, capture_output=True, text=True, timeout=600, **kw)


@pytest.mark.parametrize("case", CASES)
def test_settings_and_refvalues_bitwise_equal_python_mirror(case):
    yml, _ = C.case_paths(case)
    r = _run(SELFTEST, "settings", yml)
    assert r.returncode == 0, r.stderr
    got = dict(ln.split(" ", 1) for ln in r.stdout.splitlines())
    c = C.load_settings(yml).to_c()
    want = {
        "dim": c.dim, "particle_distance": c.particle_distance, "dt": c.dt, "cfl_condition": c.cfl_condition,
        "domain_min_x": c.domain_min[0], "domain_max_x": c.domain_max[0], "domain_min_y": c.domain_min[1],
        "domain_max_y": c.domain_max[1], "domain_min_z": c.domain_min[2], "domain_max_z": c.domain_max[2],
        "density": c.density, "kinematic_viscosity": c.kinematic_viscosity, "gravity_x": c.gravity[0],
        "gravity_y": c.gravity[1], "gravity_z": c.gravity[2], "re_pressure": c.re_pressure,
        "re_viscosity": c.re_viscosity, "re_surface": c.re_surface, "re_max": c.re_max,
        "surface_threshold": c.surface_threshold, "quasi_compressibility_on": c.quasi_compressibility_on,
        "compressibility": c.compressibility, "higher_order_source_on": c.higher_order_source_on,
        "higher_order_gamma": c.higher_order_gamma, "relaxation_coefficient": c.relaxation_coefficient,
        "collision_distance": c.collision_distance, "restitution": c.restitution,
        "n0_pressure": c.n0_pressure, "lambda_pressure": c.lambda_pressure, "n0_viscosity": c.n0_viscosity,
        "lambda_viscosity": c.lambda_viscosity, "n0_surface": c.n0_surface, "lambda_surface": c.lambda_surface,
    }
    for k, v in want.items():
        assert float(got[k]) == float(v), (k, got[k], v)  # %.17g round-trips: bit equality


def test_refvalues_pin():
    # SURVEY §8c PROBE values of the reference: n0 for ratio 3.1 in 2-D, lambda at l0 = 0.01
    yml, _ = C.case_paths("dambreak")
    got = dict(ln.split(" ", 1) for ln in _run(SELFTEST, "settings", yml).stdout.splitlines())
    assert float(got["n0_pressure"]) == 18.976416631802078
    assert float(got["lambda_pressure"]) == 2.4484722939950338e-4


@pytest.mark.parametrize("case", ["dambreak", "dambreak_fine", "hydrostatic"])
def test_csv_reader_and_writer_roundtrip_the_input_file_bytewise(case, tmp_path):
    """input.csv -> std::vector<Particle> -> Exporter::toCsv reproduces the file byte for byte (the input files
    equal the reference's shipped ones, tests/test_cases.py), so layout, float spelling and line ends match."""
    yml, csv = C.case_paths(case)
    r = _run(SELFTEST, "reexport", yml, str(tmp_path))
    assert r.returncode == 0, r.stderr
    assert open(tmp_path / "particles.csv", "rb").read() == open(csv, "rb").read()
    vtu = open(tmp_path / "particles.vtu").read().splitlines()
    n = int(r.stdout.split()[1])
    assert vtu[0] == "<?xml version='1.0' encoding='UTF-8'?>" and vtu[-1] == "</VTKFile>"
    assert f"<Piece NumberOfCells='{n}' NumberOfPoints='{n}'>" in vtu
    assert len(vtu) == 9 * n + 36  # 9 per-particle lines (position, 5 point arrays, 3 cell arrays) + 36 fixed lines


def test_float_spellings():
    r = _run(SELFTEST, "float", "gcc5", "0.3", "-1.25", "2.0", "1.99999", "123.456789", "0.000001", "0", "-0.5")
    assert r.stdout.split() == ["0.30000", "-1.25000", "2.0", "1.99999", "123.45678", "0.00000", "0.0", "-0.50000"]
    r = _run(SELFTEST, "float", "fixed6", "0.3", "-1.25", "2.0", "1e-7")
    assert r.stdout.split() == ["0.300000", "-1.250000", "2.000000", "0.000000"]


def test_fixed6_spelling_equals_printf_for_random_doubles():
    """The exporter formats with std::to_chars (fast path); the reference's clang build goes through
    std::to_string = printf("%f"). Same bytes for 4 000 doubles over 30 decades, both signs, ties included."""
    import numpy as np

    rng = np.random.default_rng(11)
    vals = np.concatenate([rng.standard_normal(2000) * 10.0 ** rng.integers(-12, 18, 2000),
                           np.round(rng.standard_normal(1000), 6) + 5e-7,  # neighbours of rounding ties
                           rng.integers(-10**6, 10**6, 1000) / 64.0])      # exactly representable fractions
    args = [repr(float(x)) for x in vals]
    r = _run(SELFTEST, "float", "fixed6", *args)
    assert r.returncode == 0
    assert r.stdout.split() == ["%f" % float(a) for a in args]


def test_gcc5_spelling_equals_the_reference_csv_writer_for_random_doubles():
    """MPSB_CSV_FLOAT=gcc5 must spell doubles as a gcc build of the reference does: csv-parser's own digit-by-digit
    routine (csv.hpp:6636-6681), here run from the reference's header itself (oracle/_ref/csv_float.exe)."""
    import numpy as np

    probe = os.path.join(ROOT, "oracle", "_ref", "csv_float.exe")
    if not os.path.exists(probe):
        pytest.skip("oracle/_ref/csv_float.exe not built (reference sources absent)")
    rng = np.random.default_rng(12)
    vals = np.concatenate([rng.standard_normal(2000) * 10.0 ** rng.integers(-8, 16, 2000),
                           np.round(rng.standard_normal(1000), 5),              # values on the 5-decimal grid
                           rng.integers(-10**7, 10**7, 500).astype(float),      # integers
                           10.0 ** rng.integers(-6, 15, 200), [0.0, -0.0, 0.99999, 0.999999, 9.999995, 1e15, 123456789012345678.0]])
    args = [repr(float(x)) for x in vals]
    ours = _run(SELFTEST, "float", "gcc5", *args)
    ref = _run(probe, *args)
    assert ours.returncode == 0 and ref.returncode == 0
    a, b = ours.stdout.split(), ref.stdout.split()
    bad = [(x, y, z) for x, y, z in zip(args, a, b) if y != z]
    assert len(a) == len(b) == len(args) and not bad, bad[:10]


def test_yaml_errors_are_fatal_and_named(tmp_path):
    bad = tmp_path / "settings.yml"
    bad.write_text("dim: 2\nparticle distance: 0.01\n")
    r = _run(SELFTEST, "settings", str(bad))
    assert r.returncode != 0 and "dt" in r.stdout + r.stderr


def test_cli_without_gpu_fails_loudly(tmp_path):
    try:
        import torch

        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    yml = tmp_path / "settings.yml"
    shutil.copy(C.case_paths("dambreak")[0], yml)
    r = _run(MPS_EXE, "-s", str(yml), "--max-steps", "1")
    assert r.returncode != 0 and "no CPU path" in r.stdout + r.stderr


def _short_case(tmp, case, finish, interval):
    S = C.load_settings(C.case_paths(case)[0])
    d = tmp / case
    d.mkdir()
    return C.write_settings_yaml(S, str(d / "settings.yml"), finish_time=finish, output_interval=interval)


def _without_pressure(text, kind):
    lines = text.splitlines()
    if kind == "csv":
        return "\n".join(",".join(ln.split(",")[:8] + ln.split(",")[9:]) if i >= 3 else ln for i, ln in enumerate(lines))
    beg = next(i for i, ln in enumerate(lines) if "Name='Pressure'" in ln)
    end = next(i for i in range(beg, len(lines)) if lines[i] == "</DataArray>")
    return "\n".join(lines[:beg] + lines[end:])


@pytest.mark.gpu
@pytest.mark.parametrize("case,finish,interval", [("dambreak", 0.06, 0.02), ("dambreak_fine", 0.0105, 0.005)])
def test_cli_outputs_match_reference_binary(case, finish, interval, tmp_path):
    """./mps -s settings.yml vs oracle/_ref/main.exe (the unmodified reference, gcc build) on the same case:
    every exported CSV and VTU byte-identical (CSV in the gcc float spelling), same step count, same
    Courant column in time_step_report.csv."""
    if not os.path.exists(REF_EXE):
        pytest.skip("oracle/_ref/main.exe not built")
    (tmp_path / "a").mkdir()
    (tmp_path / "b").mkdir()
    ya = _short_case(tmp_path / "a", case, finish, interval)
    yb = _short_case(tmp_path / "b", case, finish, interval)
    ra = _run(MPS_EXE, "-s", ya, "--quiet", env=dict(os.environ, MPSB_CSV_FLOAT="gcc5"))
    assert ra.returncode == 0, ra.stdout[-2000:] + ra.stderr[-2000:]
    rb = _run(REF_EXE, "-s", yb)
    assert rb.returncode == 0, rb.stdout[-2000:] + rb.stderr[-2000:]
    da, db = os.path.dirname(ya) + "/result", os.path.dirname(yb) + "/result"
    for sub in ("csv", "vtu"):
        fa, fb = sorted(os.listdir(f"{da}/{sub}")), sorted(os.listdir(f"{db}/{sub}"))
        assert fa == fb and len(fa) >= 3
        for f in fa:
            a, b = open(f"{da}/{sub}/{f}").read(), open(f"{db}/{sub}/{f}").read()
            if "_0000" in f:  # the reference prints an UNINITIALISED pressure at t = 0 (SURVEY B-1); ours is 0
                a, b = _without_pressure(a, sub), _without_pressure(b, sub)
            assert a == b, (sub, f)
    rep_a = [ln.split(",") for ln in open(f"{da}/time_step_report.csv").read().splitlines()]
    rep_b = [ln.split(",") for ln in open(f"{db}/time_step_report.csv").read().splitlines()]
    assert len(rep_a) == len(rep_b)
    assert rep_a[0][:12] == rep_b[0]
    for a, b in zip(rep_a[1:], rep_b[1:]):
        for col in (0, 1, 2, 3, 4, 10, 11):  # step, dt, times, finish, files written, Courant
            assert a[col] == b[col], (col, a, b)


def read_state(path):
    """Binary state of mps_b200/host/state_io.hpp -> dict of arrays."""
    import numpy as np

    raw = open(path, "rb").read()
    assert raw[:8] == b"MPSBST1\x00"
    n = int(np.frombuffer(raw, np.int64, 1, 8)[0])
    out = {"time": float(np.frombuffer(raw, np.float64, 1, 16)[0]), "n": n}
    off = 24
    for name, dt, cnt in (("type", np.int32, n), ("boundary_condition", np.int32, n), ("position", np.float64, 3 * n),
                          ("velocity", np.float64, 3 * n), ("pressure", np.float64, n), ("number_density_ratio", np.float64, n)):
        out[name] = np.frombuffer(raw, dt, cnt, off)
        off += cnt * np.dtype(dt).itemsize
    assert off == len(raw)
    out["position"] = out["position"].reshape(n, 3)
    out["velocity"] = out["velocity"].reshape(n, 3)
    return out


@pytest.mark.parametrize("case,nx", [("dambreak", 0), ("dambreak_fine", 0), ("hydrostatic", 0), ("dambreak_fine", 400)])
def test_cpp_generator_equals_python_generator_bitwise(case, nx, tmp_path):
    """host/generator.hpp (C++) against mps_b200.cases.generate_lattice (the generator the golden fixtures and the
    bench use): same ids, types and full-precision positions; for the shipped sizes also the reference's input.csv."""
    import numpy as np

    out = tmp_path / "s.bin"
    args = [GEN_EXE, "--case", case, "--state", str(out), "--csv", str(tmp_path / "input.csv")]
    if nx:
        args += ["--nx", str(nx)]
    r = _run(*args)
    assert r.returncode == 0, r.stderr
    st = read_state(out)
    if nx:
        S, t, p, v = C.scaled_dambreak(nx, base=case)
    else:
        g = C.GEOMETRIES[case]
        t, p = C.generate_lattice(g["l0"], g["domain_x"], g["domain_y"], g["fluid_x"], g["fluid_y"])
        assert open(tmp_path / "input.csv", "rb").read() == open(C.case_paths(case)[1], "rb").read()
    assert st["n"] == len(t) and np.array_equal(st["type"], t) and np.array_equal(st["position"], p)
    assert not st["velocity"].any() and st["time"] == 0.0


@pytest.mark.gpu
def test_checkpoint_restart_is_bitwise_an_uninterrupted_run(tmp_path):
    """mps --state-out after k steps, then --state-in for m more steps == k + m steps straight: every field of the
    final binary states identical (higher-order source term on, so the saved ratio and flags matter)."""
    import numpy as np

    S = C.load_settings(C.case_paths("dambreak_fine_higher_order")[0])
    dt = S.dt

    def run(tag, finish, state_in=None):
        d = tmp_path / tag
        d.mkdir()
        yml = C.write_settings_yaml(S, str(d / "settings.yml"), finish_time=finish, output_interval=1.0)
        out = d / "final.bin"
        cmd = [MPS_EXE, "-s", yml, "--quiet", "--state-out", str(out)] + (["--state-in", str(state_in)] if state_in else [])
        r = _run(*cmd)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        return read_state(out)

    straight = run("straight", 30.5 * dt)
    first = run("first", 12.5 * dt)
    second = run("second", 30.5 * dt, tmp_path / "first" / "final.bin")
    assert abs(first["time"] - 12 * dt) < 1e-12 and abs(second["time"] - straight["time"]) < 1e-15
    for k in ("type", "boundary_condition", "position", "velocity", "pressure", "number_density_ratio"):
        assert np.array_equal(second[k], straight[k]), k


@pytest.mark.parametrize("n", [0, 1, 1000])
def test_binary_state_round_trips_bitwise_and_refuses_damaged_files(n, tmp_path):
    """host/state_io.hpp on the CPU: denormals, -0.0, 0.1 + 0.2 and 1e300 come back bit for bit; a truncated file
    and a CSV under the same name are refused."""
    r = _run(SELFTEST, "state", str(n), str(tmp_path / "state.bin"))
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.split() == ["particles", str(n), "bytes", str(24 + 72 * n), "mismatches", "0",
                                "truncated_refused", "1", "text_refused", "1"]


def test_cpp_rendezvous_shares_the_id_between_three_processes():
    """host/rendezvous.hpp: rank 0 serves a 128-byte id over TCP on MASTER_PORT + 1, the others fetch it."""
    port = 30100 + os.getpid() % 500
    procs = []
    for rank in (1, 2, 0):  # clients first: they must retry until the server is up
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="3", LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([SELFTEST, "rendezvous", "x"], stdout=subprocess.PIPE, text=True, env=env))
    outs = [p.communicate(timeout=120)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    sums = {o.split()[-1] for o in outs}
    assert len(sums) == 1 and sums != {"0"}, outs
