"""Spatial decomposition of one replica over several ranks (csrc/gc_decomp.cuh): the
trajectory must be BIT-identical to the undecomposed one for every world size -- each
bead's force is a gather over its full neighbourhood and the noise is keyed by (seed,
bead, step).  The GPU test emulates the ranks with several contexts on one device
(LocalExchange); the NCCL path is exercised by scripts/run_decomp.py under torchrun."""
import numpy as np
import pytest

from gcpkg import gc_b200
from helpers import small_interphase, stages

decomp = gc_b200.decomp


def test_slot_ranges_partition():
    for n, w in [(10, 1), (10, 3), (60672, 8), (7, 8), (2_000_000, 8)]:
        r = decomp.slot_ranges(n, w)
        assert r[0][0] == 0 and r[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
        assert max(hi - lo for lo, hi in r) == (n + w - 1) // w


def test_all_gather_layout_gloo_world2(tmp_path):
    """The exchange the ranks perform (all_gather_into_tensor of equal chunks) places
    rank k's buffer at [k * elems, (k + 1) * elems) -- what k_decomp_scatter reads."""
    import torch.multiprocessing as mp
    mp.spawn(_gloo_worker, args=(str(tmp_path / "rdv"),), nprocs=2, join=True)


def _gloo_worker(rank, rdv):
    import torch
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"file://{rdv}", rank=rank, world_size=2)
    elems = 5
    send = torch.full((elems,), float(rank + 1))
    recv = torch.zeros(2 * elems)
    dist.all_gather_into_tensor(recv, send)
    assert recv.tolist() == [1.0] * elems + [2.0] * elems
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 3])
def test_decomposed_trajectory_is_bit_identical(world):
    d, cfg, x = small_interphase()
    steps = 40

    def make():
        g = gc_b200.System(d.n, 1, 0)
        stages.setup_interphase(g, d, cfg, contacts=False)
        g.set_positions(x)
        g.set_seeds([77])
        stages.interphase_callback0(g, cfg)
        return g

    ref = make()
    ref.run(steps, cfg.timestep, restart=True)
    xr, cr = ref.get_positions(0), ref.get_context(0)
    ref.close()

    ranks = [make() for _ in range(world)]
    ex = decomp.LocalExchange(ranks)
    ex.begin(cfg.timestep)
    ex.step(steps)
    for g in ranks:
        np.testing.assert_array_equal(g.get_positions(0), xr)
        c = g.get_context(0)
        assert c.time == cr.time and list(c.semiaxes) == list(cr.semiaxes)
        assert c.core_scale == cr.core_scale
        g.close()


@pytest.mark.gpu
def test_decomposed_context_rejects_plain_run():
    d, cfg, x = small_interphase()
    g = gc_b200.System(d.n, 1, 0)
    stages.setup_interphase(g, d, cfg, contacts=False)
    g.set_positions(x)
    g.decomp_configure(0, 2)
    with pytest.raises(gc_b200.api.GcError):
        g.run(5, cfg.timestep)
    g.close()
