"""Host-side inputs: settings, RefValues, lattice generator (CPU)."""
import os

import numpy as np
import pytest

from mps_b200 import cases as C

REF = "/root/reference"


def test_ref_values_pins():
    # SURVEY §8c pins (measured from the compiled reference)
    n0, lam = C.ref_values(2, 0.01, 3.1 * 0.01)
    assert n0 == 18.976416631802078
    assert lam == 2.4484722939950338e-4
    assert C.ref_values(2, 0.005, 3.1 * 0.005)[1] == 6.1211807349875844e-5
    assert C.ref_values(2, 0.004, 3.1 * 0.004)[1] == 3.9175556703920528e-5


@pytest.mark.parametrize("case,n,counts", [("dambreak", 458, (200, 82, 176)), ("dambreak_fine", 6146, (5000, 378, 768)),
                                           ("hydrostatic", 2958, (2400, 182, 376))])
def test_generator_counts_and_csv_roundtrip(case, n, counts, tmp_path):
    t, p = C.generate_lattice(**C.GEOMETRIES[case])
    assert len(t) == n
    assert ((t == C.FLUID).sum(), (t == C.WALL).sum(), (t == C.DUMMY_WALL).sum()) == counts
    path = tmp_path / "input.csv"
    C.write_input_csv(str(path), t, p)
    with open(path) as f, open(C.case_paths(case)[1]) as g:
        assert f.read() == g.read()  # the committed input.csv is what the generator writes
    time, t2, p2, v2 = C.load_input_csv(str(path))
    assert time == 0.0 and np.array_equal(t, t2) and np.allclose(p, p2, atol=5e-7) and not v2.any()


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not present")
@pytest.mark.parametrize("case", ["dambreak", "dambreak_fine", "hydrostatic"])
def test_input_csv_identical_to_reference(case):
    with open(C.case_paths(case)[1], "rb") as f, open(f"{REF}/inputs/{case}/input.csv", "rb") as g:
        assert f.read() == g.read()


def test_settings_parse_and_struct():
    s = C.load_settings(C.case_paths("dambreak_fine")[0])
    assert s.dim == 2 and s.particle_distance == 0.004 and s.dt == 0.0005
    assert s.quasi_compressibility_on and s.compressibility == 0.45e-09 and not s.higher_order_source_on
    assert s.re_max == 3.1 * 0.004 and s.collision_distance == 0.5 * 0.004
    c = s.to_c()
    assert c.n0_pressure == 18.976416631802078 and c.domain_min[0] == -0.1 and c.domain_max[1] == 0.5
    h = C.load_settings(C.case_paths("dambreak_fine_higher_order")[0])
    assert h.higher_order_source_on and h.higher_order_gamma == 1.0e-03
    o = C.load_settings(C.case_paths("dambreak_fine_original")[0])
    assert not o.quasi_compressibility_on and o.relaxation_coefficient == 1.0


def test_scaled_dambreak_counts():
    s, t, p, v = C.scaled_dambreak(200)  # l0 = 0.002, the survey's "generated l0=.002" case
    assert len(t) == 22268 and (t == C.FLUID).sum() == 200 * 100
    assert abs(s.dt / s.particle_distance - 0.125) < 1e-12
