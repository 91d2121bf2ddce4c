"""Generates tests/golden/reference_vectors.npz from the REFERENCE's own code.

Runs oracle/_ref/libgc_ref_strict.so -- micromd headers and the reference's
contact_map.cpp / kinetochore_fiber_forcefield.cpp compiled unmodified from
/root/reference, strict IEEE flags -- on small seeded inputs and records inputs and
outputs.  /root/reference does not exist on the GPU box, so the vectors are committed;
tests/test_golden.py checks the port oracle (CPU) and the CUDA path (GPU) against them.

    python scripts/make_golden.py          (needs /root/reference; ~10 s)
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "oracle"))
sys.path.insert(0, str(ROOT / "tests"))
import oracle_py as O  # noqa: E402
from gcpkg import gc_b200  # noqa: E402
from golden_cases import CASES, build_case  # noqa: E402

KIND = "strict"


def main():
    O.build(("strict",))
    out = {"oracle_kind": np.asarray(KIND),
           "reference_commit": np.asarray("snsinfu/3d-genome-cycle 9069a11, micromd a886d68")}
    for name in CASES:
        res = build_case(name, lambda n: O.OracleSystem(n, KIND), is_oracle=True)
        for k, v in res.items():
            out[f"{name}/{k}"] = np.asarray(v)
        print(name, {k: np.asarray(v).shape for k, v in res.items()})
    path = ROOT / "tests" / "golden" / "reference_vectors.npz"
    np.savez_compressed(path, **out)
    print("wrote", path, path.stat().st_size, "bytes")


if __name__ == "__main__":
    main()
