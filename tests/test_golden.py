"""Committed reference outputs (tests/golden/reference_vectors.npz, produced by
scripts/make_golden.py from the reference's own code compiled in place) against
  * the port oracle          -- CPU, tight tolerances: pins the oracle that travels;
  * the CUDA path (C-ABI)    -- GPU, the north-star bars: pair/contact sets bit-exact,
    forces and positions <= 1e-5 relative (FP32 device vs FP64 reference)."""
from pathlib import Path

import numpy as np
import pytest

import oracle_py as O
from gcpkg import gc_b200
from golden_cases import CASES, build_case
from helpers import rel_force_err

GOLD = np.load(Path(__file__).resolve().parent / "golden" / "reference_vectors.npz")


def compare(name, got, tol_force, tol_energy, tol_pos, tol_dt, exact_traj_contacts):
    keys = [k.split("/", 1)[1] for k in GOLD.files if k.startswith(name + "/")]
    assert keys
    for k in keys:
        want, have = GOLD[f"{name}/{k}"], np.asarray(got[k])
        if k == "x":
            np.testing.assert_array_equal(have, want)          # the case builder is unchanged
        elif k.startswith("force_"):
            if np.abs(want).max() == 0:
                assert np.abs(have).max() == 0, k
            else:
                assert rel_force_err(have, want) < tol_force, k
        elif k.startswith("energy_"):
            assert float(have) == pytest.approx(float(want), rel=tol_energy, abs=1e-6 * tol_energy / 1e-11), k
        elif k.startswith("x_"):
            assert np.abs(have - want).max() < tol_pos, k
        elif k == "ctx":
            np.testing.assert_allclose(have, want, rtol=1e-7 if tol_pos > 1e-9 else 1e-12)
        elif k == "axial_reaction":
            assert np.abs(want).max() > 0
            assert np.abs(have - want).max() < tol_force * np.abs(want).max()
        elif k == "reaction_force":
            assert float(have) == pytest.approx(float(want), rel=tol_force)
        elif k == "dt":
            assert float(have) == pytest.approx(float(want), rel=tol_dt)
        elif k == "contacts" and not exact_traj_contacts:
            a = {(i, j): v for i, j, v in have.tolist()}
            b = {(i, j): v for i, j, v in want.tolist()}
            diff = sum(abs(a.get(q, 0) - b.get(q, 0)) for q in set(a) | set(b))
            assert diff <= max(2, int(2e-3 * sum(b.values())))
        else:                                                  # pairs_*, triples, contacts
            np.testing.assert_array_equal(have, want, err_msg=k)


def test_golden_file_is_from_the_reference():
    assert str(GOLD["oracle_kind"]) == "strict"
    assert "micromd a886d68" in str(GOLD["reference_commit"])
    assert sorted({k.split("/")[0] for k in GOLD.files if "/" in k}) == sorted(CASES)


@pytest.mark.parametrize("name", CASES)
def test_port_oracle_matches_reference_vectors(name):
    O.build(("port",))
    got = build_case(name, lambda n: O.OracleSystem(n, "port"), is_oracle=True)
    compare(name, got, tol_force=1e-11, tol_energy=1e-11, tol_pos=1e-10, tol_dt=1e-11,
            exact_traj_contacts=True)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_path_matches_reference_vectors(name):
    got = build_case(name, lambda n: gc_b200.System(n, 1, 0), is_oracle=False)
    compare(name, got, tol_force=1e-5, tol_energy=2e-5, tol_pos=1e-5 * 2.0, tol_dt=1e-5,
            exact_traj_contacts=False)
