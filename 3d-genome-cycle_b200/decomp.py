"""One large polymer spread over several GPUs (BASELINE configs[4]): host side of the
slot-range decomposition in csrc/gc_decomp.cuh.  One process per GPU (torchrun); every
rank holds all beads, evaluates and moves its slab, and ONE NCCL all-gather per step
exchanges the new positions.  `LocalExchange` emulates the collective between several
contexts on one device (tests on a single GPU)."""
from __future__ import annotations


def slot_ranges(n_beads: int, world: int):
    """[lo, hi) of sorted cell-list slots owned by each rank (gc_decomp_configure)."""
    chunk = (n_beads + world - 1) // world
    return [(min(r * chunk, n_beads), min(r * chunk + chunk, n_beads)) for r in range(world)]


class DecomposedRun:
    """Drives gc_decomp_step_compute -> all-gather -> gc_decomp_step_commit."""

    def __init__(self, system, rank: int, world: int, device: int = 0, group=None):
        import torch
        self.sys, self.rank, self.world, self.group = system, rank, world, group
        system.decomp_configure(rank, world)
        self.send = self.recv = None
        if world > 1:
            self.send, self.recv = system.decomp_buffers(device)
        self.stream = torch.cuda.ExternalStream(system.stream(), device=device)

    def begin(self, timestep, temperature=1.0):
        """Temperature / timestep, restart, first grid plan (gc_run with 0 steps)."""
        self.sys.run(0, timestep, temperature=temperature, restart=True)

    def step(self, steps=1):
        import torch
        import torch.distributed as dist
        s = self.sys
        if self.world == 1:
            raise RuntimeError("world = 1: use System.run")
        with torch.cuda.stream(self.stream):
            for _ in range(steps):
                s.decomp_step_compute()
                dist.all_gather_into_tensor(self.recv, self.send, group=self.group)
                s.decomp_step_commit()


class LocalExchange:
    """`world` contexts on ONE device standing in for `world` ranks: the all-gather is
    a set of device-to-device copies.  Used by the single-GPU parity test."""

    def __init__(self, systems, device=0):
        import torch
        self.systems = systems
        self.world = len(systems)
        for r, s in enumerate(systems):
            s.decomp_configure(r, self.world)
        self.bufs = [s.decomp_buffers(device) for s in systems]
        self.elems = self.bufs[0][0].numel()
        self.torch = torch

    def begin(self, timestep, temperature=1.0):
        for s in self.systems:
            s.run(0, timestep, temperature=temperature, restart=True)

    def step(self, steps=1):
        for _ in range(steps):
            for s in self.systems:
                s.decomp_step_compute()
            for s in self.systems:
                s.synchronize()
            for _, recv in self.bufs:
                for k, (send, _) in enumerate(self.bufs):
                    recv[k * self.elems:(k + 1) * self.elems].copy_(send)
            self.torch.cuda.synchronize()
            for s in self.systems:
                s.decomp_step_commit()
