"""CPU check of the large-N fixtures (tests/golden/large_nx*.npz, recorded from the UNMODIFIED compiled
reference by tests/golden/make_large_golden.py): the inputs regenerate bit for bit from (nx, seed) on this
machine, and the oracle restatement (oracle/mps_oracle.cpp) reproduces the reference's digests, iteration
count and sampled solution at the sizes it finishes in seconds. The GPU side of the same fixtures is
tests/test_gpu_parity.py::test_large_generated_case_against_reference."""
import numpy as np
import pytest

from mps_b200 import cases as C
from oracle import refbind as R
from tests import util as U


def test_fixtures_cover_the_bench_size():
    have = {(int(np.load(p)["nx"]), int(np.load(p)["jitter"])) for p in U.large_golden_files()}
    assert {(400, 0), (400, 1), (800, 0), (1416, 0)} <= have
    for p in U.large_golden_files():
        z = np.load(p)
        assert int(z["cg_info"]) == 0 and float(z["cg_error"]) <= 2.3e-16
        assert len(z["sample_ids"]) == len(z["sample_solve"]) > 1000


@pytest.mark.parametrize("tag", ["400", "400j"])
def test_oracle_port_reproduces_the_reference_fixture(tag):
    path = [p for p in U.large_golden_files() if p.endswith(f"large_nx{tag}.npz")][0]
    z = np.load(path)
    S, t, p, v = U.large_case(z)
    orc = R.OracleMPS(S.to_c(), t, p, v)
    for ph in (R.PH_NEIGHBORS, R.PH_GRAVITY, R.PH_VISCOSITY, R.PH_MOVE, R.PH_NEIGHBORS, R.PH_COLLISION, R.PH_NEIGHBORS):
        orc.phase(ph)
    assert U.sha(*orc.neighbors()) == str(z["nbr_digest"])
    for ph in (R.PH_BC, R.PH_SOURCE, R.PH_MATRIX, R.PH_DIRICHLET):
        orc.phase(ph)
    assert U.sha(orc.get_state(["boundary_condition"])["boundary_condition"]) == str(z["bc_digest"])
    assert U.sha(*orc.ppe()) == str(z["ppe_digest"])
    x, it, err, info = orc.cg_solve()
    assert it == int(z["cg_iterations"]) and err == float(z["cg_error"])
    assert np.array_equal(x[z["sample_ids"]], z["sample_solve"])  # the port follows Eigen's operation order
