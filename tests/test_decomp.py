"""Spatial decomposition of one replica over several ranks (csrc/gc_decomp.cuh): z-slabs of
the cell grid with a halo exchange between neighbouring ranks.  The trajectory must be
BIT-identical to the undecomposed one for every world size -- each bead's force is a gather
over its full neighbourhood and the noise is keyed by (seed, bead, step, epoch).  The GPU
tests emulate the ranks with several contexts on one device (LocalExchange); the NCCL path
is exercised by scripts/run_decomp.py and bench.py under torchrun."""
import numpy as np
import pytest

from gcpkg import gc_b200
from helpers import small_interphase, stages

decomp = gc_b200.decomp
synth = gc_b200.synth


def test_slab_starts_partition_with_equal_volumes():
    for world, layers in [(1, 10), (2, 12), (3, 30), (8, 66), (8, 130), (8, 11)]:
        z = decomp.slab_starts(world, layers)
        assert z[0] == 1 and z[-1] == layers - 1
        assert all(a <= b for a, b in zip(z, z[1:]))
    # volume of the unit ball between the cuts: equal to within one layer's worth
    layers = 1002
    z = np.asarray(decomp.slab_starts(8, layers), dtype=np.float64)
    u = 2.0 * (z - 1.0) / (layers - 2) - 1.0
    vol = np.diff((2.0 + 3.0 * u - u ** 3) / 4.0)
    assert np.allclose(vol, 1.0 / 8, atol=3.0 / layers)


def test_exchange_pattern_gloo_world2(tmp_path):
    """The collectives a step performs, on CPU tensors over gloo: header all-gather places
    rank k's header at [16 k, 16 (k + 1)); send_hi of rank 0 lands in recv_lo of rank 1 and
    send_lo of rank 1 in recv_hi of rank 0."""
    import torch.multiprocessing as mp
    mp.spawn(_gloo_worker, args=(str(tmp_path / "rdv"),), nprocs=2, join=True)


def _gloo_worker(rank, rdv):
    import torch
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"file://{rdv}", rank=rank, world_size=2)
    H = decomp.HEADER_DOUBLES
    hdr = torch.full((H,), float(rank + 1), dtype=torch.float64)
    allh = torch.zeros(2 * H, dtype=torch.float64)
    dist.all_gather_into_tensor(allh, hdr)
    assert allh.tolist() == [1.0] * H + [2.0] * H
    elems = 8
    send_lo, send_hi = torch.full((elems,), 10.0 + rank), torch.full((elems,), 20.0 + rank)
    recv_lo, recv_hi = torch.zeros(elems), torch.zeros(elems)
    ops = []
    if rank == 0:
        ops = [dist.P2POp(dist.isend, send_hi, 1), dist.P2POp(dist.irecv, recv_hi, 1)]
    else:
        ops = [dist.P2POp(dist.isend, send_lo, 0), dist.P2POp(dist.irecv, recv_lo, 0)]
    for w in dist.batch_isend_irecv(ops):
        w.wait()
    if rank == 0:
        assert recv_hi.tolist() == [11.0] * elems and recv_lo.tolist() == [0.0] * elems
    else:
        assert recv_lo.tolist() == [20.0] * elems and recv_hi.tolist() == [0.0] * elems
    dist.destroy_process_group()


def _interphase_rank(d, cfg, x, seed=77):
    g = gc_b200.System(d.n, 1, 0)
    stages.setup_interphase(g, d, cfg, contacts=False)
    g.set_positions(x)
    g.set_seeds([seed])
    stages.interphase_callback0(g, cfg)
    return g


@pytest.mark.gpu
@pytest.mark.parametrize("peer", [False, True])
def test_decomposed_trajectory_is_bit_identical(peer):
    """A whole (scaled-down) nucleus on two ranks: chains, nucleolar bonds, the nucleolar
    droplet (whose targets may not be current on a rank: they are then beyond its reach), the
    wall ODE.  The grid has about ten layers, so two slabs of five with a three-layer halo
    is all it affords; larger worlds run on the polymer below."""
    world = 2
    d, cfg, x = small_interphase()
    steps = 40
    ref = _interphase_rank(d, cfg, x)
    ref.run(steps, cfg.timestep, restart=True)
    xr, cr = ref.get_positions(0), ref.get_context(0)
    ref.close()

    ranks = [_interphase_rank(d, cfg, x) for _ in range(world)]
    ex = decomp.LocalExchange(ranks, halo_layers=3, peer=peer)
    ex.begin(cfg.timestep)
    ex.step(steps)
    np.testing.assert_array_equal(ex.gather_positions(), xr)
    for g in ranks:
        c = g.get_context(0)
        assert c.time == cr.time and list(c.semiaxes) == list(cr.semiaxes)
        assert c.core_scale == cr.core_scale and c.reaction_force == cr.reaction_force
        g.close()


def _polymer_rank(d, x, cfg):
    g = gc_b200.System(d.n, 1, 0)
    stages.setup_interphase(g, d, cfg, contacts=False)
    g.set_positions(x)
    g.set_seeds([5])
    stages.interphase_callback0(g, cfg)
    return g


@pytest.mark.gpu
@pytest.mark.parametrize("world,peer", [(2, False), (3, True), (8, False), (8, True)])
def test_decomposed_200k_polymer_is_bit_identical_and_exchanges_a_surface(world, peer):
    """peer = True: the pack kernel writes straight into the neighbours' receive buffers (on one
    device here, over NVLink peer memory between processes) under the old grid + one layer.
    BASELINE configs[4] at a size the suite affords: 2 x 10^5 beads, 30 steps.  Bit-equal
    positions, every bead owned exactly once, balanced slabs, and a halo that is a surface
    term (the round-1 scheme shipped every bead to every rank)."""
    n, steps = 200_000, 30
    semi = tuple(s * (n / 1e5) ** (1 / 3) for s in (5.0, 4.0, 3.0))
    cfg = synth.InterphaseConfig(wall_semiaxes_init=semi, nucleolus_droplet_energy=0.0,
                                 wall_semiaxes_spring=(3e4 * n / 60672,) * 3)
    d, x = synth.make_polymer(n, seed=4, semiaxes=semi)
    ref = _polymer_rank(d, x, cfg)
    ref.run(steps, cfg.timestep, restart=True)
    xr = ref.get_positions(0)
    ref.close()

    ranks = [_polymer_rank(d, x, cfg) for _ in range(world)]
    # the walk's bonds (0.3) span two cell edges (0.15): a three-layer position halo
    ex = decomp.LocalExchange(ranks, halo_layers=3, peer=peer)
    ex.begin(cfg.timestep)
    ex.step(steps // 2)
    # resizing the halo buffers mid-run keeps the halos just received
    cap = int(1.5 * max(max(c) for c in ex.halo_counts())) + 1024
    ex.set_halo(3, cap, cfg.timestep)
    ex.step(steps - steps // 2)
    np.testing.assert_array_equal(ex.gather_positions(), xr)
    owned = [int(g.decomp_owned().sum()) for g in ranks]
    assert sum(owned) == n
    sent = [lo + hi for lo, hi in ex.halo_counts()]
    # equal-volume slabs: balanced ownership; the halo is a surface term -- a small fraction
    # of a thick slab (2 ranks: ~25 layers each), most of a thin one (8 ranks: ~6 layers)
    # (8 ranks: a layer is a sixth of a slab here, and inner slabs keep a minimum thickness)
    assert max(owned) < (1.25 if world < 8 else 1.5) * n / world, owned
    if world == 8:
        print("owned per rank after", steps, "steps:", owned)
    if world == 2:
        assert max(sent) < 0.3 * max(owned), (sent, owned)
    for g in ranks:
        g.close()


@pytest.mark.gpu
def test_decomposed_context_rejects_plain_run_and_reports_a_thin_halo():
    d, cfg, x = small_interphase()
    g = _interphase_rank(d, cfg, x)
    g.decomp_configure(0, 2)
    with pytest.raises(gc_b200.api.GcError):
        g.run(5, cfg.timestep)
    g.close()
    # 16 ranks over a grid of about ten layers: inner slabs thinner than the halo
    ranks = [_interphase_rank(d, cfg, x) for _ in range(16)]
    ex = decomp.LocalExchange(ranks)
    with pytest.raises(gc_b200.api.GcError, match="halo"):
        ex.begin(cfg.timestep)
        ex.step(1)
        for r in ranks:
            r.get_positions(0)
    for r in ranks:
        r.close()


@pytest.mark.gpu
def test_decomposed_run_resumes_from_a_checkpoint():
    """(positions merged from the ranks, context, step) saved mid-run and loaded into FRESH
    decomposed contexts continues bit for bit: gc_set_step re-stamps every bead as current."""
    n, world = 60_000, 3
    semi = tuple(s * (n / 1e5) ** (1 / 3) for s in (5.0, 4.0, 3.0))
    cfg = synth.InterphaseConfig(wall_semiaxes_init=semi, nucleolus_droplet_energy=0.0,
                                 wall_semiaxes_spring=(3e4 * n / 60672,) * 3)
    d, x = synth.make_polymer(n, seed=4, semiaxes=semi)
    ref = _polymer_rank(d, x, cfg)
    ref.run(40, cfg.timestep, restart=True)
    xr = ref.get_positions(0)
    ref.close()

    ranks = [_polymer_rank(d, x, cfg) for _ in range(world)]
    ex = decomp.LocalExchange(ranks, halo_layers=3, peer=True)
    ex.begin(cfg.timestep)
    ex.step(25)
    xa, ca = ex.gather_positions(), ranks[0].get_context(0)
    for g in ranks:
        g.close()

    ranks = [_polymer_rank(d, xa, cfg) for _ in range(world)]
    ex = decomp.LocalExchange(ranks, halo_layers=3, peer=True)
    ex.begin(cfg.timestep)                                   # restart: epoch 1, as in the first run
    for g in ranks:
        g.set_context(time=ca.time, semiaxes=list(ca.semiaxes), core_scale=ca.core_scale,
                      bond_scale=ca.bond_scale, contact_distance=ca.contact_distance)
        g.set_step(25)
        g.run(0, cfg.timestep, restart=False)                # the grid follows the loaded context
    ex.step(15)
    np.testing.assert_array_equal(ex.gather_positions(), xr)
    assert ranks[1].get_context(0).step == 40
    for g in ranks:
        g.close()
