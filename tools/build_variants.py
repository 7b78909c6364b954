"""Builds experiment copies of libmps_b200.so under build/exp/ with -D overrides of the PCG kernel's compile-time
switches (csrc/k_pcg.cu: PCG_NWARPS, PCG_NSTAGES, PCG_NCH, PCG_NBATCH, PCG_B_DESC, PCG_C_DESC, PCG_STREAM_HINT,
PCG_VEC_HINT). tools/pcg_variants.sh then times the solve of each on the GPU box.
    python tools/build_variants.py l2          # the L2-reuse set behind profiles/r1_v23_pcg_l2_reuse_variants.txt
    python tools/build_variants.py name=-DPCG_NWARPS=16,-DPCG_NCH=16 ..."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mps_b200 import build as B  # noqa: E402

SETS = {
    "l2": {
        "v0": ["-DPCG_C_DESC=0"],
        "v1_Bd": ["-DPCG_C_DESC=0", "-DPCG_B_DESC=1"],
        "v2_Cd": ["-DPCG_C_DESC=1"],
        "v3_BdCd": ["-DPCG_B_DESC=1", "-DPCG_C_DESC=1"],
        "v1e": ["-DPCG_C_DESC=0", "-DPCG_B_DESC=1", "-DPCG_STREAM_HINT=1"],
        "v1ev": ["-DPCG_C_DESC=0", "-DPCG_B_DESC=1", "-DPCG_STREAM_HINT=1", "-DPCG_VEC_HINT=1"],
        "v0ev": ["-DPCG_C_DESC=0", "-DPCG_STREAM_HINT=1", "-DPCG_VEC_HINT=1"],
    },
}

if __name__ == "__main__":
    variants = {}
    for a in sys.argv[1:]:
        if a in SETS:
            variants.update(SETS[a])
        elif "=" in a:
            name, flags = a.split("=", 1)
            variants[name] = flags.split(",")
        else:
            raise SystemExit(f"unknown set {a!r}; known: {sorted(SETS)}")
    out = os.path.join(B.ROOT, "build", "exp")
    os.makedirs(out, exist_ok=True)
    for name, flags in variants.items():
        B.build_library(out=os.path.join(out, f"lib_{name}.so"), extra=flags)
        print(name, " ".join(flags))
