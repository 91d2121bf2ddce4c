"""One large polymer spread over several GPUs (BASELINE configs[4]): host side of the
slab + halo decomposition in csrc/gc_decomp.cuh.  One process per GPU (torchrun); every
rank owns a z-slab of the cell grid and keeps current copies of the beads within a few
cell layers of it.  Per step: one all-gather of 128-byte headers (global bounding box,
wall sums, error flag) and one send/recv of halo beads with each of the two neighbouring
ranks -- surface traffic, not the whole system.  `LocalExchange` emulates the exchange
between several contexts on one device (tests on a single GPU)."""
from __future__ import annotations

import numpy as np

HEADER_DOUBLES = 16


def slab_starts(world: int, layers: int, margin: int = 1):
    """First grid layer of every rank's slab plus the end of the last one (mirror of
    slab_start in csrc/gc_kernels.cuh): equal-volume cuts of the ellipsoid inscribed in the
    `layers - 2 margin` interior layers.  The volume fraction of an ellipsoid below height
    u in [-1, 1] is (2 + 3u - u^3) / 4; u(f) = 2 cos((acos(1 - 2f) + 4 pi) / 3) inverts it."""
    nint = layers - 2 * margin
    out = []
    for k in range(world + 1):
        if k <= 0:
            out.append(margin)
        elif k >= world:
            out.append(margin + nint)
        else:
            f = k / world
            u = 2.0 * np.cos((np.arccos(1.0 - 2.0 * f) + 4.0 * np.pi) / 3.0)
            z = margin + int(np.floor(0.5 * (u + 1.0) * nint + 0.5))
            out.append(min(max(z, margin), margin + nint))
    return out


class DecomposedRun:
    """Drives gc_decomp_step_compute -> header all-gather -> gc_decomp_step_pack ->
    neighbour send/recv -> gc_decomp_step_commit on the context's own stream."""

    @property
    def scheme(self):
        how = ("packed straight into the neighbours' buffers over NVLink peer memory (CUDA IPC); one header "
               "all-gather per step" if self.peer else "header all-gather + NCCL send/recv with the two neighbours")
        return "z-slabs of the cell grid + halo exchange: " + how

    def __init__(self, system, rank: int, world: int, device: int = 0, group=None,
                 halo_layers: int | None = None, halo_capacity: int = 0, graph_steps: int = 0,
                 peer: bool = False):
        import torch
        self.sys, self.rank, self.world, self.group = system, rank, world, group
        self.device = device
        system.decomp_configure(rank, world)
        self.buf = None
        self.graph_steps = int(graph_steps)
        self._graph = None
        self._eager_done = False
        self.halo_layers = halo_layers or 2
        if world > 1:
            if halo_layers is not None or halo_capacity:
                system.decomp_set_halo(self.halo_layers, halo_capacity)
            self.buf = system.decomp_buffers(device)
        self.peer = bool(peer) and world > 1
        if self.peer:
            self._attach_peers()
        self.stream = torch.cuda.ExternalStream(system.stream(), device=device)

    def _attach_peers(self):
        """Peer mode: every rank maps its neighbours' receive buffers (CUDA IPC) and packs the
        halo straight into them over NVLink; the header all-gather is then the only
        collective of a step."""
        import torch.distributed as dist
        blobs = [None] * self.world
        dist.all_gather_object(blobs, self.sys.decomp_peer_export(), group=self.group)
        import torch
        err = None
        try:
            self.sys.decomp_peer_attach_ipc(blobs[self.rank - 1] if self.rank > 0 else None,
                                            blobs[self.rank + 1] if self.rank + 1 < self.world else None)
        except Exception as ex:             # e.g. no peer access between two devices
            err = ex
        ok = torch.tensor([0 if err else 1], device=f"cuda:{self.device}")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)      # also the barrier
        if int(ok.item()) == 0:             # every rank takes the same way out
            self.sys.decomp_peer_detach()
            raise RuntimeError(f"peer mode unavailable on at least one rank ({err})")

    def _detach_peers(self):
        import torch.distributed as dist
        self.sys.decomp_peer_detach()
        dist.barrier(group=self.group)          # nobody resizes before everybody has let go

    def begin(self, timestep, temperature=1.0):
        """Temperature / timestep, restart, first grid plan (gc_run with 0 steps); every rank
        must have been given all N positions."""
        self._run_args = (timestep, temperature)
        self.sys.run(0, timestep, temperature=temperature, restart=True)

    def tune_capacity(self, slack=1.5, floor=1024):
        """Shrinks the halo buffers (whole buffers travel every step) to `slack` x the largest
        halo any rank shipped in the last step, and the per-bead kernel's grid to `slack` x the
        largest bead count any rank owns; call after a few warm-up steps.  Collective."""
        import torch
        import torch.distributed as dist
        if self.world == 1:
            return 0
        self.sys.synchronize()
        t = torch.tensor([max(self.halo_counts()), int(self.sys.decomp_owned().sum())],
                         device=f"cuda:{self.device}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        cap = int(slack * int(t[0].item())) + floor
        self.invalidate_graph()
        if self.peer:
            self._detach_peers()
        self.sys.decomp_set_halo(self.halo_layers, cap)
        if self.peer:
            self._attach_peers()
        self.sys.decomp_set_owned_bound(int(slack * int(t[1].item())) + floor)
        timestep, temperature = self._run_args
        self.sys.run(0, timestep, temperature=temperature, restart=False)   # parameters only
        self.buf = self.sys.decomp_buffers(self.device)
        self._graph = None
        return cap

    @property
    def exchange_bytes_per_step(self):
        """Bytes this rank sends per step: its header to every rank plus the two halo buffers
        (whole buffers travel; the bead count rides in their first element)."""
        if self.world == 1:
            return 0
        sides = (self.rank > 0) + (self.rank + 1 < self.world)
        if self.peer:                       # only the packed beads cross NVLink
            halo = 16 * sum(self.halo_counts()) + 16 * sides
            return HEADER_DOUBLES * 8 * (self.world - 1) + halo
        halo = self.buf["send_lo"].numel() * 4
        return HEADER_DOUBLES * 8 * (self.world - 1) + sides * halo

    def halo_counts(self):
        """(beads sent down, beads sent up) in the last step."""
        lo = int(self.buf["send_lo"][:1].view(dtype=_i32()).item())
        hi = int(self.buf["send_hi"][:1].view(dtype=_i32()).item())
        return lo, hi

    def step(self, steps=1):
        """Advances `steps` steps.  With graph_steps > 0 (constructor), whole blocks of that
        many steps -- kernels and NCCL operations alike -- are replayed from one CUDA graph
        captured on first use; the remainder runs eagerly."""
        import torch
        if self.world == 1:
            raise RuntimeError("world = 1: use System.run")
        K = self.graph_steps
        with torch.cuda.stream(self.stream):
            while K and steps >= K and self._eager_done:
                if self._graph is None:
                    self._capture(K)
                self._graph.replay()
                self.sys.decomp_host_advance(K)
                steps -= K
            self._eager(steps)

    def _capture(self, K):
        import torch
        self.sys.synchronize()              # (eager steps came first: the communicators are live)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=self.stream, capture_error_mode="thread_local"):
            self._eager(K)
        self.sys.decomp_host_advance(-K)    # the capture moved the host mirror, not the device
        self._graph = g

    def invalidate_graph(self):
        """After anything that changes buffers or parameters (tune_capacity, set_context ...);
        also before the process group is destroyed (a live graph holds NCCL kernels)."""
        import torch
        if self._graph is not None:
            torch.cuda.synchronize()
            self._graph = None

    def _eager(self, steps):
        import torch.distributed as dist
        s, b = self.sys, self.buf
        ops_template = []
        if self.rank + 1 < self.world:
            ops_template += [(dist.isend, b["send_hi"], self.rank + 1), (dist.irecv, b["recv_hi"], self.rank + 1)]
        if self.rank > 0:
            ops_template += [(dist.isend, b["send_lo"], self.rank - 1), (dist.irecv, b["recv_lo"], self.rank - 1)]
        if steps > 0:
            self._eager_done = True
        for _ in range(steps):
            s.decomp_step_compute()
            dist.all_gather_into_tensor(b["headers_all"], b["header"], group=self.group)
            s.decomp_step_pack()
            if not self.peer:
                ops = [dist.P2POp(op, t, self._peer(peer), group=self.group) for op, t, peer in ops_template]
                for w in dist.batch_isend_irecv(ops):
                    w.wait()
            s.decomp_step_commit()

    def close(self):
        """Before the context or the process group goes away."""
        self.invalidate_graph()
        if self.peer:
            self._detach_peers()
            self.peer = False

    def timed_phases(self, steps=20):
        """Eager steps with CUDA events between the phases: mean ms of (compute, header
        all-gather, pack, halo send/recv, commit) on this rank.  The collectives' figures
        include the wait for the slowest rank."""
        import torch
        import torch.distributed as dist
        s, b = self.sys, self.buf
        names = ("compute", "header_all_gather", "pack", "halo_send_recv", "commit")
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(6)] for _ in range(steps)]
        with torch.cuda.stream(self.stream):
            for k in range(steps):
                ev[k][0].record(self.stream)
                s.decomp_step_compute()
                ev[k][1].record(self.stream)
                dist.all_gather_into_tensor(b["headers_all"], b["header"], group=self.group)
                ev[k][2].record(self.stream)
                s.decomp_step_pack()
                ev[k][3].record(self.stream)
                ops = []
                if self.rank + 1 < self.world:
                    ops += [dist.P2POp(dist.isend, b["send_hi"], self._peer(self.rank + 1), group=self.group),
                            dist.P2POp(dist.irecv, b["recv_hi"], self._peer(self.rank + 1), group=self.group)]
                if self.rank > 0:
                    ops += [dist.P2POp(dist.isend, b["send_lo"], self._peer(self.rank - 1), group=self.group),
                            dist.P2POp(dist.irecv, b["recv_lo"], self._peer(self.rank - 1), group=self.group)]
                if not self.peer:
                    for w in dist.batch_isend_irecv(ops):
                        w.wait()
                ev[k][4].record(self.stream)
                s.decomp_step_commit()
                ev[k][5].record(self.stream)
        s.synchronize()
        return {n: sum(e[i].elapsed_time(e[i + 1]) for e in ev) / steps for i, n in enumerate(names)}

    def _peer(self, rank):
        import torch.distributed as dist
        if self.group is None:
            return rank
        return dist.get_global_rank(self.group, rank)

    def gather_positions(self):
        """All N positions on every rank (float64 [N, 3]): each rank contributes the beads it
        owns.  Host-side (N x 24 bytes through the process group) -- for checks and output,
        not part of a step."""
        import torch
        import torch.distributed as dist
        own = self.sys.decomp_owned()
        x = self.sys.get_positions(0)
        x[~own] = 0.0
        t = torch.from_numpy(np.concatenate([x, own[:, None].astype(np.float64)], axis=1)).cuda(self.device)
        dist.all_reduce(t, group=self.group)
        out = t.cpu().numpy()
        if not np.all(out[:, 3] == 1.0):
            raise RuntimeError("decomposition: %d beads are not owned exactly once"
                               % int(np.sum(out[:, 3] != 1.0)))
        return out[:, :3].copy()


def _i32():
    import torch
    return torch.int32


class LocalExchange:
    """`world` contexts on ONE device standing in for `world` ranks: the collectives are
    device-to-device copies.  Used by the single-GPU parity tests."""

    def __init__(self, systems, device=0, halo_layers=None, halo_capacity=0, peer=False):
        import torch
        self.systems = systems
        self.world = len(systems)
        for r, s in enumerate(systems):
            s.decomp_configure(r, self.world)
            if halo_layers is not None or halo_capacity:
                s.decomp_set_halo(halo_layers or 2, halo_capacity)
        self.device = device
        self.bufs = [s.decomp_buffers(device) for s in systems]
        self.torch = torch
        self.peer = bool(peer)
        if self.peer:
            self._attach()

    def _attach(self):
        recv = [s.decomp_recv_buffers() for s in self.systems]
        for k, s in enumerate(self.systems):
            s.decomp_peer_attach(recv[k - 1] if k > 0 else None, recv[k + 1] if k + 1 < self.world else None)

    def begin(self, timestep, temperature=1.0):
        for s in self.systems:
            s.run(0, timestep, temperature=temperature, restart=True)

    def step(self, steps=1):
        H = HEADER_DOUBLES
        for _ in range(steps):
            for s in self.systems:
                s.decomp_step_compute()
            for s in self.systems:
                s.synchronize()
            for b in self.bufs:
                for k, src in enumerate(self.bufs):
                    b["headers_all"][k * H:(k + 1) * H].copy_(src["header"])
            self.torch.cuda.synchronize()
            for s in self.systems:
                s.decomp_step_pack()
            for s in self.systems:
                s.synchronize()
            if not self.peer:
                for k in range(self.world - 1):
                    self.bufs[k + 1]["recv_lo"].copy_(self.bufs[k]["send_hi"])
                    self.bufs[k]["recv_hi"].copy_(self.bufs[k + 1]["send_lo"])
                self.torch.cuda.synchronize()
            for s in self.systems:
                s.decomp_step_commit()

    def set_halo(self, layers, capacity, timestep, temperature=1.0):
        """Resizes every rank's halo buffers between two steps."""
        for s in self.systems:
            s.synchronize()
            if self.peer:
                s.decomp_peer_detach()
        for s in self.systems:
            s.decomp_set_halo(layers, capacity)
        if self.peer:
            self._attach()
        for s in self.systems:
            s.run(0, timestep, temperature=temperature, restart=False)   # parameters only
        self.bufs = [s.decomp_buffers(self.device) for s in self.systems]

    def halo_counts(self):
        out = []
        for b in self.bufs:
            out.append((int(b["send_lo"][:1].view(dtype=self.torch.int32).item()),
                        int(b["send_hi"][:1].view(dtype=self.torch.int32).item())))
        return out

    def gather_positions(self):
        """float64 [N, 3] merged from the beads each emulated rank owns."""
        n = self.systems[0].n
        x = np.zeros((n, 3))
        owners = np.zeros(n, dtype=np.int32)
        for s in self.systems:
            own = s.decomp_owned()
            x[own] = s.get_positions(0)[own]
            owners += own
        if not np.all(owners == 1):
            raise RuntimeError("decomposition: %d beads are not owned exactly once" % int(np.sum(owners != 1)))
        return x
