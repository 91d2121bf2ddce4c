"""Ensemble sharding and the cross-rank reductions (host logic of the multi-GPU path).
CPU: world_size-2 gloo processes.  GPU: the device-side reduce/merge behind them."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from gcpkg import gc_b200  # noqa: E402

E = gc_b200.ensemble


def test_shard_replicas_partition():
    for total, world in [(64, 8), (64, 3), (5, 8), (1, 1), (0, 2)]:
        ids = [list(E.shard_replicas(total, world, r)) for r in range(world)]
        assert sorted(sum(ids, [])) == list(range(total))
        sizes = [len(i) for i in ids]
        assert max(sizes) - min(sizes) <= 1


def test_replica_seeds_do_not_depend_on_sharding():
    a = E.replica_seeds(1000, range(64))
    b = np.concatenate([E.replica_seeds(1000, E.shard_replicas(64, 8, r)) for r in range(8)])
    np.testing.assert_array_equal(a, b)
    assert len(set(a.tolist())) == 64
    assert not np.array_equal(a, E.replica_seeds(1001, range(64)))


def _synthetic_map(seed, n_beads=500, n_pairs=4000):
    rng = np.random.default_rng(seed)
    i = rng.integers(0, n_beads - 1, n_pairs)
    j = i + 1 + rng.integers(0, 20, n_pairs)
    key = (i.astype(np.int64) << 32) | np.minimum(j, n_beads - 1).astype(np.int64)
    key, cnt = np.unique(key, return_counts=True)
    return torch.from_numpy(key), torch.from_numpy(cnt.astype(np.int32) * (seed + 1))


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        keys, counts = _synthetic_map(rank, n_pairs=3000 + 1500 * rank)    # ragged sizes
        k, c = E.gather_contacts(keys, counts)
        obs = E.reduce_sum(torch.tensor([float(rank + 1), 1.0], dtype=torch.float64))
        empty = E.gather_contacts(torch.empty(0, dtype=torch.int64),
                                  torch.empty(0, dtype=torch.int32))
        torch.save((k, c, obs, empty), f"{out}/rank{rank}.pt")
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_gather_contacts_gloo_world2(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(tmp_path / f"rank{r}.pt") for r in range(world)]
    # expected: plain sum of the two synthetic maps
    want = {}
    for r in range(world):
        k, c = _synthetic_map(r, n_pairs=3000 + 1500 * r)
        for kk, cc in zip(k.tolist(), c.tolist()):
            want[kk] = want.get(kk, 0) + cc
    wk = np.asarray(sorted(want))
    wc = np.asarray([want[q] for q in wk])
    for k, c, obs, empty in res:
        np.testing.assert_array_equal(k.numpy(), wk)              # sorted, merged, on every rank
        np.testing.assert_array_equal(c.numpy(), wc)
        assert obs.tolist() == [3.0, 2.0]
        assert empty[0].numel() == 0 and empty[1].numel() == 0
    ijv = E.keys_to_ijv(res[0][0], res[0][1])
    assert ijv.dtype == np.uint32 and (ijv[:, 0] < ijv[:, 1]).all()
    assert (np.diff(ijv[:, 0].astype(np.int64) * 2 ** 32 + ijv[:, 1]) > 0).all()


def test_contact_probability_known_answer():
    chains = np.asarray([(0, 5), (5, 8)])
    # every (i, i+1) pair of chain 0 seen twice, one (i, i+2) pair once, one trans pair
    ijv = np.asarray([(0, 1, 2), (1, 2, 2), (2, 3, 2), (3, 4, 2), (0, 2, 1), (4, 5, 9), (5, 7, 4)],
                     dtype=np.uint32)
    p = E.contact_probability(ijv, chains, n_updates=2, n_cells=1)
    # s = 1: counts 8 over (4 + 2) available pairs x 2 updates; s = 2: (1 + 4) / ((3 + 1) x 2)
    assert p[1] == pytest.approx(8 / 12) and p[2] == pytest.approx(5 / 8) and p[0] == 0


@pytest.mark.gpu
def test_device_reduce_and_merge_match_host_fetch():
    from helpers import build_pair, small_interphase
    d, cfg, x = small_interphase(scale=0.02, radius=0.6)
    R = 3
    g = gc_b200.System(d.n, R, 0)
    gc_b200.stages.setup_interphase(g, d, cfg)
    rng = np.random.default_rng(4)
    for r in range(R):
        g.set_positions(x + rng.normal(size=x.shape) * 0.01, r)
    g.set_context(contact_distance=0.12)
    g.contacts_update()
    host = g.contacts_fetch(-1)                                    # ensemble sum through the host path
    k, c = g.contacts_reduce_device(-1)
    np.testing.assert_array_equal(E.keys_to_ijv(k, c), host)
    assert c.max().item() >= 2                                     # replicas do share contacts
    # merging the per-replica runs must give the same ensemble map, and equal plain torch
    parts = [g.contacts_reduce_device(r) for r in range(R)]
    kk, cc = torch.cat([p[0] for p in parts]), torch.cat([p[1] for p in parts])
    mk, mc = g.contacts_merge_device(kk, cc)
    np.testing.assert_array_equal(E.keys_to_ijv(mk, mc), host)
    tk, tc = E.merge_contacts_torch(kk, cc)
    assert torch.equal(tk, mk) and torch.equal(tc, mc)
    ek, ec = g.contacts_merge_device(kk[:0], cc[:0])
    assert ek.numel() == 0 and ec.numel() == 0
