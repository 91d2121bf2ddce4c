"""Ensembles of independent cell trajectories across the GPUs of one box.

The reference has no distributed layer: users launch `scripts/simulate -s <seed>` once
per cell and `src/cool.py:80-114` merges the per-file contact pixels afterwards
(SURVEY.md 2.2).  Here the cells of an ensemble are sharded over ranks (one process per
GPU, no data-path collective while stepping) and the contact maps and observables are
reduced over NCCL at output cadence.  Everything in this module is host logic over
`torch.distributed`, so it runs unchanged on CPU tensors with the gloo backend (tests).
"""
from __future__ import annotations

import numpy as np


def shard_replicas(n_total: int, world: int, rank: int) -> range:
    """Contiguous block of global cell ids owned by `rank` (sizes differ by at most 1)."""
    base, extra = divmod(n_total, world)
    start = rank * base + min(rank, extra)
    return range(start, start + base + (1 if rank < extra else 0))


def replica_seeds(master_seed: int, cell_ids) -> np.ndarray:
    """One 64-bit Philox key per cell: splitmix64 of (master_seed + cell id), so a cell's
    trajectory does not depend on how the ensemble is sharded."""
    out = []
    for c in cell_ids:
        z = (int(master_seed) + 0x9E3779B97F4A7C15 * (int(c) + 1)) & 0xFFFFFFFFFFFFFFFF
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
        out.append(z ^ (z >> 31))
    return np.asarray(out, dtype=np.uint64)


def replica_seeds_reference(master_base: int, cell_ids, stage: str = "interphase") -> np.ndarray:
    """Seeds as a reference ensemble derives them: cell c is `scripts/simulate -s (master_base +
    c)`, whose `prepare` turns the master seed into per-stage seeds with std::seed_seq and whose
    stage driver draws the seed of every simulate call from std::mt19937_64(stage seed)
    (prepare.py: run_seeds; SIM/stage_prepare/prepare.cpp:549-562).  The value returned for a
    cell is the one the reference hands md::simulate_brownian_dynamics for `stage`."""
    from .prepare import run_seeds
    return np.asarray([run_seeds(int(master_base) + int(c))[stage] for c in cell_ids], dtype=np.uint64)


def merge_contacts_torch(keys, counts):
    """Reference merge (plain torch, any device): sort + sum of equal keys."""
    import torch
    if keys.numel() == 0:
        return keys, counts
    uk, inv = torch.unique(keys, sorted=True, return_inverse=True)
    out = torch.zeros(uk.numel(), dtype=torch.int64, device=keys.device)
    out.index_add_(0, inv, counts.to(torch.int64))
    return uk, out.to(counts.dtype)


def gather_contacts(keys, counts, merge=merge_contacts_torch, group=None):
    """Ensemble contact map: every rank contributes its sorted (key, count) runs, all
    ranks receive the merged map.  Variable lengths are handled by an all_gather of the
    sizes followed by an all_gather of padded buffers.  `merge` is the product's device
    merge (System.contacts_merge_device) on GPUs and plain torch in the CPU tests."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return keys, counts
    world = dist.get_world_size(group)
    n = torch.tensor([keys.numel()], dtype=torch.int64, device=keys.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    nmax = max(max(sizes), 1)
    kpad = torch.zeros(nmax, dtype=keys.dtype, device=keys.device)
    cpad = torch.zeros(nmax, dtype=counts.dtype, device=keys.device)
    kpad[:keys.numel()] = keys
    cpad[:counts.numel()] = counts
    kall = [torch.empty_like(kpad) for _ in range(world)]
    call = [torch.empty_like(cpad) for _ in range(world)]
    dist.all_gather(kall, kpad, group=group)
    dist.all_gather(call, cpad, group=group)
    k = torch.cat([kall[r][:sizes[r]] for r in range(world)])
    c = torch.cat([call[r][:sizes[r]] for r in range(world)])
    return merge(k, c)


def reduce_sum(values, group=None):
    """All-reduce (sum) of a small tensor of observables (Rg^2 sums, energies, counts)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(values, op=dist.ReduceOp.SUM, group=group)
    return values


_PINNED = None


def keys_to_ijv(keys, counts) -> np.ndarray:
    """(i << 32 | j, count) -> the (i, j, count) triples contact_map::accumulate returns
    (SIM/stage_interphase/contact_map.cpp:66-85).  CUDA tensors are unpacked on the device
    and come back in one copy through pinned memory."""
    import torch
    k = keys.detach()
    out = torch.empty((k.numel(), 3), dtype=torch.int32, device=k.device)
    out[:, 0] = (k >> 32).to(torch.int32)
    out[:, 1] = k.to(torch.int32)                      # low word (two's-complement wrap)
    out[:, 2] = counts.detach().to(torch.int32)
    if out.is_cuda:
        global _PINNED
        if _PINNED is None or _PINNED.numel() < out.numel():     # page-locking is slow: keep it
            _PINNED = torch.empty(int(out.numel() * 1.25) + 1024, dtype=torch.int32, pin_memory=True)
        host = _PINNED[:out.numel()].view(out.shape)
        host.copy_(out)
        return host.numpy().view(np.uint32)          # valid until the next call
    return out.numpy().view(np.uint32)


def contact_probability(ijv: np.ndarray, chains: np.ndarray, n_updates: int, n_cells: int,
                        max_s: int | None = None) -> np.ndarray:
    """P(s): mean contact frequency of intra-chain bead pairs at genomic separation s
    (the cis decay profile of src/pc1.py:156-183, on raw counts): sum of counts of
    pairs with j - i = s on the same chain / (pairs available x updates x cells)."""
    chains = np.asarray(chains)
    n = int(chains[:, 1].max())
    chain_of = np.full(n + 1, -1, dtype=np.int64)
    for c, (s, e) in enumerate(chains):
        chain_of[s:e] = c
    lengths = chains[:, 1] - chains[:, 0]
    max_s = int(max_s or lengths.max() - 1)
    i, j, v = ijv[:, 0].astype(np.int64), ijv[:, 1].astype(np.int64), ijv[:, 2].astype(np.float64)
    ok = (i <= n) & (j <= n)
    i, j, v = i[ok], j[ok], v[ok]
    cis = (chain_of[i] == chain_of[j]) & (chain_of[i] >= 0)
    sep = (j - i)[cis]
    hist = np.bincount(sep[sep <= max_s], weights=v[cis][sep <= max_s], minlength=max_s + 1)
    avail = np.zeros(max_s + 1)
    for L in lengths:
        s = np.arange(1, min(L - 1, max_s) + 1)
        avail[s] += L - s
    with np.errstate(divide="ignore", invalid="ignore"):
        p = np.where(avail > 0, hist / (avail * max(n_updates, 1) * max(n_cells, 1)), 0.0)
    p[0] = 0.0
    return p
