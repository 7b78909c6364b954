"""The C-ABI library builds, loads and exports every symbol include/mps_b200.h declares (CPU)."""
import ctypes as ct
import os
import re

import numpy as np
import pytest

from mps_b200 import build, capi, cases as C
from tests import util as U


@pytest.fixture(scope="module")
def lib():
    build.build_library()
    return capi.load_library()


def test_header_symbols_exported(lib):
    with open(os.path.join(U.ROOT, "include", "mps_b200.h")) as f:
        src = f.read()
    declared = set(re.findall(r"\b(mpsb_[a-z_0-9]+)\s*\(", src))
    assert declared == set(capi.EXPORTED)
    for name in declared:
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.mpsb_version()


def test_struct_sizes_match_header(tmp_path):
    # compile the header with gcc and compare sizeof with the ctypes images
    import subprocess

    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "mps_b200.h"\nint main(){printf("%zu %zu %zu\\n",'
                   'sizeof(mpsb_settings),sizeof(mpsb_step_stats),sizeof(mpsb_fields));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(U.ROOT, "include"), str(src), "-o", str(exe)], check=True)
    a, b, c = map(int, subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split())
    assert (a, b, c) == (ct.sizeof(C.CSettings), ct.sizeof(capi.CStepStats), ct.sizeof(capi.CFields))


def test_no_cpu_fallback(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    S, t, p, v = C.load_case("dambreak")
    with pytest.raises(capi.MpsError) as e:
        capi.DeviceMPS(S, t, p, v)
    assert e.value.code == -2 and "no CPU path" in str(e.value)


def test_bad_arguments(lib):
    h = ct.c_void_p()
    assert lib.mpsb_create(None, 0, None, None, None, 0, ct.byref(h)) == -1
    assert lib.mpsb_step(None, 0, None) == -1
    assert lib.mpsb_num_particles(None) == 0


def test_every_environment_switch_the_library_reads_is_documented_in_the_header():
    """include/mps_b200.h lists the MPSB_* switches; a getenv() in csrc/ or host/ that the header does not mention fails here."""
    import glob
    import os
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "mps_b200.h")).read()
    read = set()
    for path in glob.glob(os.path.join(root, "mps_b200", "csrc", "*.cu*")) + glob.glob(os.path.join(root, "mps_b200", "host", "*.[ch]pp")):
        read |= set(re.findall(r'getenv\("(MPSB_[A-Z0-9_]+)"\)', open(path).read()))
    assert read, "no switches found: the pattern is stale"
    missing = sorted(v for v in read if v not in header)
    assert not missing, missing
