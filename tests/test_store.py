"""Output path (store.py): the reference's on-disk schema as a flat directory tree."""
import json

import numpy as np
import pytest

import oracle_py as O
from gcpkg import gc_b200

synth, P, S = gc_b200.synth, gc_b200.prepare, gc_b200.store


class _Ctx:
    time, semiaxes, core_scale, bond_scale = 0.125, (2.0, 2.5, 3.0), 0.75, 0.625


def test_frames_contexts_contacts_round_trip(tmp_path, oracle_kind):
    st = S.FlatStore(tmp_path / "traj")
    st.set_stage("interphase")
    assert st.load_steps() == [] and not st.check_positions(1000)
    rng = np.random.default_rng(1)
    x = rng.normal(size=(500, 3)) * 2.0
    st.save_positions(1000, x, quantize=True)             # host data: the store rounds like the reference
    st.save_interphase_context(1000, _Ctx(), mean_energy=5.5)
    st.append_frame(1000)
    st.save_positions(2000, O.quantize_f32(x, 16, oracle_kind))   # device frames arrive rounded
    st.append_frame(2000)
    ijv = np.array([[0, 1, 3], [0, 7, 1], [4, 9, 2]], dtype=np.uint32)
    st.save_contacts(2000, ijv)
    st.save_contacts(3000, np.zeros((0, 3), dtype=np.uint32))     # empty: nothing is written
    assert st.load_steps() == [1000, 2000]
    assert json.loads((tmp_path / "traj/stages/interphase/.steps.json").read_text()) == ["1000", "2000"]
    want = O.quantize_f32(x, 16, oracle_kind)
    np.testing.assert_array_equal(st.load_positions(1000), want.astype(np.float64))
    np.testing.assert_array_equal(st.load_positions(2000), want.astype(np.float64))
    assert np.load(tmp_path / "traj/stages/interphase/1000/positions.npy").dtype == np.float32
    c = st.load_interphase_context(1000)
    assert list(c) == ["time", "wall_semiaxes", "core_scale", "bond_scale", "mean_energy", "wall_energy"]
    assert c["wall_semiaxes"] == [2.0, 2.5, 3.0] and c["mean_energy"] == 5.5
    got = st.load_contacts(2000)
    assert got.dtype == np.int32 and got.shape == (3, 3)
    np.testing.assert_array_equal(got, ijv.astype(np.int32))
    assert not (tmp_path / "traj/stages/interphase/3000").exists()
    with pytest.raises(S.StoreError):
        st.load_positions(4000)
    st.clear_frames()
    assert st.load_steps() == []
    # another stage has its own frame index
    st.set_stage("prometaphase")
    st.append_frame(5)
    assert st.load_steps() == [5]
    with pytest.raises(S.StoreError):
        S.H5Store(tmp_path / "x.h5") if not _has_h5py() else (_ for _ in ()).throw(S.StoreError("skip"))


def _has_h5py():
    try:
        import h5py  # noqa: F401
        return True
    except ImportError:
        return False


def test_prepare_metadata_layout(tmp_path):
    d = synth.make_diploid_design(scale=1.0)
    chains = P.chains_from_design(d)
    p = P.prepare(chains, master_seed=1000)
    st = S.FlatStore(tmp_path / "traj")
    st.write_prepared(p, config_json="{}", chains_source=P.format_chains_tsv(chains)[:200], master_seed=1000)
    assert st.load_seed("interphase") == P.stage_seeds(1000)["interphase"]
    assert st.load_seed("relaxation") == st.load_seed("interphase")          # linked metadata
    assert st.load_seed("telophase") == st.load_seed("anaphase") == P.stage_seeds(1000)["anaphase"]
    np.testing.assert_array_equal(st.load_chains("interphase"), d.chains)
    ab = st._read_array("/stages/interphase/metadata/ab_factors")
    assert ab.dtype == np.float32 and ab.shape == (d.n, 2)
    types = st._read_array("/stages/interphase/metadata/particle_types")
    assert (types == S.INTERPHASE_TYPES["nucleolus"]).sum() == 30 and (types == S.INTERPHASE_TYPES["active_nor"]).sum() == 15
    assert st._read_array("/stages/prometaphase/metadata/pole_positions").shape == (2, 3)
    assert st._read_array("/stages/prometaphase/metadata/sister_chromatids").shape == (46, 2)
    assert len(st._read_strings("/stages/prometaphase/metadata/chain_names")) == 92
    # a second handle on the same directory sees the links
    st2 = S.FlatStore(tmp_path / "traj", create=False)
    assert st2.load_seed("relaxation") == st.load_seed("interphase")


@pytest.mark.gpu
def test_device_frames_are_the_store_rounding(oracle_kind, tmp_path):
    """gc_get_frame / gc_frame_begin+end == quantize() of simulation_store.cpp:22-32 (restated in
    oracle/ref_harness.cpp) applied to the positions; frames flow into the store through the
    cycle runner's frame sink."""
    d = synth.make_diploid_design(scale=0.05)
    cfg = synth.InterphaseConfig()
    g = gc_b200.System(d.n, 2, 0)
    gc_b200.stages.setup_interphase(g, d, cfg)
    x = synth.globule_positions(d, seed=2, radius=0.75)
    g.set_positions(x)
    gc_b200.stages.begin_interphase(g, cfg)
    g.run(100, cfg.timestep)
    for bits in (16, 12, 20):
        f = g.get_frame(-1, bits)
        for r in range(2):
            np.testing.assert_array_equal(f[r], O.quantize_f32(g.get_positions(r), bits, oracle_kind))
    g.frame_begin(1, 16)
    np.testing.assert_array_equal(g.frame_end()[0], O.quantize_f32(g.get_positions(1), 16, oracle_kind))
    g.close()
    st = S.FlatStore(tmp_path / "cycle")
    w = S.TrajectoryWriter(st, sampling_interval=1000)
    rep = gc_b200.cycle.run_cell_cycle(design=d, scale_steps=3e-3, seed=3, frame_sink=w)
    saved = 0
    for stage in ("anaphase", "telophase", "interphase", "prometaphase"):
        st.set_stage(stage)
        saved += len(st.load_steps())
    assert rep["frames_saved"] == saved >= 3
    st.set_stage("interphase")
    last = st.load_steps()[-1]
    assert st.load_positions(last).shape == (d.n, 3) and np.isfinite(st.load_positions(last)).all()
    assert st.load_interphase_context(last)["core_scale"] > 0.5
