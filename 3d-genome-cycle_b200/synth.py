"""Synthetic inputs of the shapes BASELINE.json names.

The reference ships no input data (its example/ directory is git-ignored and
absent), so chains, A/B factors and start structures are generated here with
fixed seeds.  Topology rules follow the reference's prepare stage:
  SIM/stage_prepare/prepare.cpp:168-238 (chains, nucleolar beads and bonds),
  279-313 (coarse-grained mitotic chains, kinetochore bead),
  SIM/common/simulation_store.cpp:22-32 (16-bit mantissa quantisation of
  stored positions).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

# hg38 chromosome lengths in bp (harness data, not from the reference)
HG38 = {
    "chr1": 248956422, "chr2": 242193529, "chr3": 198295559, "chr4": 190214555,
    "chr5": 181538259, "chr6": 170805979, "chr7": 159345973, "chr8": 145138636,
    "chr9": 138394717, "chr10": 133797422, "chr11": 135086622, "chr12": 133275309,
    "chr13": 114364328, "chr14": 107043718, "chr15": 101991189, "chr16": 90338345,
    "chr17": 83257441, "chr18": 80373285, "chr19": 58617616, "chr20": 64444167,
    "chr21": 46709983, "chr22": 50818468, "chrX": 156040895,
}
NOR_CHROMS = ("chr13", "chr14", "chr15", "chr21", "chr22")


@dataclass
class InterphaseConfig:
    """Defaults of interphase_config (SIM/common/simulation_config.hpp:61-115)."""
    temperature: float = 1.0
    timestep: float = 1e-5
    steps: int = 700000
    sampling_interval: int = 1000
    logging_interval: int = 1000
    relaxation_spacestep: float = 0.001
    relaxation_steps: int = 10000
    contactmap_distance: float = 0.24
    contactmap_update_interval: int = 20
    contactmap_output_window: int = 10
    a_core_diameter: float = 0.30
    b_core_diameter: float = 0.24
    a_core_repulsion: float = 2.5
    b_core_repulsion: float = 2.5
    a_core_bond_spring: float = 100.0
    b_core_bond_spring: float = 50.0
    a_core_bond_length: float = 0.0
    b_core_bond_length: float = 0.0
    # not in the reference's JSON traits list -> always 0
    a_core_2nd_bond_spring: float = 0.0
    b_core_2nd_bond_spring: float = 0.0
    a_core_mobility: float = 1.0
    b_core_mobility: float = 1.0
    core_scale_init: float = 0.5
    core_scale_tau: float = 0.5
    bond_scale_init: float = 0.5
    bond_scale_tau: float = 0.5
    nucleolus_bead_count: int = 2
    nucleolus_ab_factor: tuple = (0.0, 10.0)
    nucleolus_bond_spring: float = 10.0
    nucleolus_bond_length: float = 0.0
    nucleolus_droplet_energy: float = 0.3
    nucleolus_droplet_decay: float = 0.2
    nucleolus_droplet_cutoff: float = 0.4
    nucleolus_mobility: float = 1.0
    wall_semiaxes_init: tuple = (2.0, 2.0, 2.0)
    wall_semiaxes_spring: tuple = (3e4, 3e4, 3e4)
    wall_packing_spring: float = 1000.0
    wall_ab_factor: tuple = (0.0, 10.0)
    wall_mobility: float = 2e-4


@dataclass
class MitoticConfig:
    """Defaults of mitotic_phase_config (SIM/common/simulation_config.hpp:15-58)."""
    temperature: float = 1.0
    timestep: float = 1e-4
    anaphase_steps: int = 200000
    telophase_steps: int = 50000
    prometaphase_steps: int = 400000
    sampling_interval: int = 1000
    logging_interval: int = 10000
    anaphase_start_stddev: float = 1.0
    coarse_graining: int = 100
    core_diameter: float = 0.3
    core_repulsion: float = 2.0
    bond_length: float = 0.3
    bond_spring: float = 1000.0
    bending_energy: float = 1.0
    penalize_centromere_bending: bool = False
    core_mobility: float = 0.1
    sister_separation: float = 0.3
    spindle_axis: tuple = (0.0, 5.0, 0.0)
    kfiber_decay_rate_prometaphase: float = 1.0
    kfiber_decay_rate_anaphase: float = 1.0
    kfiber_length_prometaphase: float = 0.0
    kfiber_length_anaphase: float = 0.0
    polar_ejection_force: float = 0.0
    polar_ejection_cross_section: float = 0.0
    anaphase_spindle_shift: tuple = (0.0, 2.0, 0.0)
    telophase_packing_radius: float = 1.5
    telophase_packing_spring: float = 100.0
    telophase_bond_spring_multiplier: float = 1.0
    telophase_bending_energy_multiplier: float = 1.0


@dataclass
class InterphaseDesign:
    n: int
    chains: np.ndarray            # [C, 2] start, end
    names: list
    a: np.ndarray                 # [n]
    b: np.ndarray
    mobility: np.ndarray
    nucleolar_bonds: np.ndarray   # [B, 2] (nor, nucleolar bead)
    centromeres: np.ndarray       # [C, 2] centromeric range per chain
    seed: int = 0


@dataclass
class MitoticDesign:
    n: int
    chains: np.ndarray            # [C, 2]
    kinetochores: np.ndarray      # [C] bead index
    sisters: np.ndarray = field(default_factory=lambda: np.zeros((0, 2), dtype=np.int32))
    poles: np.ndarray = field(default_factory=lambda: np.zeros((2, 3)))
    seed: int = 0


def quantize_f32(x, bits=16):
    """save_positions rounding: SIM/common/simulation_store.cpp:22-32,192-216."""
    x = np.asarray(x, dtype=np.float64)
    mant, exp = np.frexp(x)
    q = np.ldexp(np.rint(np.ldexp(mant, bits)), exp - bits)
    return q.astype(np.float32)


def make_diploid_design(seed=20250001, bead_kb=100, scale=1.0, cfg: InterphaseConfig | None = None):
    """46 chains chr{1..22,X}:{a,b}; A/B/u blocks from a 2-state Markov chain
    (mean block 20 beads, 10 % u); one 30-bead centromere block at 40 % of the
    chain; 3 active NOR beads on the short arm of chr13/14/15/21/22 copy a,
    each carrying nucleolus_bead_count nucleolar beads appended after all
    chains (prepare.cpp:221-238).  `scale` < 1 shrinks every chain (tests)."""
    cfg = cfg or InterphaseConfig()
    rng = np.random.default_rng(seed)
    a_list, b_list, chains, names, cens, nors = [], [], [], [], [], []
    pos = 0
    for name, bp in HG38.items():
        for copy in ("a", "b"):
            length = max(int(math.ceil(bp / (bead_kb * 1000) * scale)), 8)
            state = int(rng.integers(0, 2))
            a = np.zeros(length)
            b = np.zeros(length)
            i = 0
            while i < length:
                blk = int(rng.geometric(1.0 / 20.0))
                u = rng.random() < 0.10
                for k in range(i, min(i + blk, length)):
                    if u:
                        a[k], b[k] = 0.5, 0.5        # src/annotate.py:138-142 'u'
                    elif state == 0:
                        a[k], b[k] = 1.0, 0.0        # 'A'
                    else:
                        a[k], b[k] = 0.0, 1.0        # 'B'
                i += blk
                state ^= 1
            cen_len = max(min(30, length // 4), 1)
            cen_start = int(0.4 * length)
            cens.append((pos + cen_start, pos + min(cen_start + cen_len, length)))
            if name in NOR_CHROMS and copy == "a":
                for k in range(3):
                    nors.append(pos + min(2 + 2 * k, length - 1))
            chains.append((pos, pos + length))
            names.append(f"{name}:{copy}")
            a_list.append(a)
            b_list.append(b)
            pos += length
    a = np.concatenate(a_list)
    b = np.concatenate(b_list)
    bonds = []
    for nor in nors:
        for _ in range(cfg.nucleolus_bead_count):
            bonds.append((nor, len(a)))
            a = np.append(a, cfg.nucleolus_ab_factor[0])
            b = np.append(b, cfg.nucleolus_ab_factor[1])
    n = len(a)
    chains = np.asarray(chains, dtype=np.int32)
    # simulation_driver_particles.cpp:19-34
    mob = np.where(a >= b, cfg.a_core_mobility, cfg.b_core_mobility).astype(np.float64)
    bonds = np.asarray(bonds, dtype=np.int32).reshape(-1, 2)
    if len(bonds):
        mob[bonds[:, 1]] = cfg.nucleolus_mobility
    return InterphaseDesign(n=n, chains=chains, names=names, a=a, b=b, mobility=mob,
                            nucleolar_bonds=bonds, centromeres=np.asarray(cens, dtype=np.int32),
                            seed=seed)


def confined_walk(n, rng, step, semiaxes, start=None):
    """Random walk of n points with fixed step length kept inside an ellipsoid."""
    sx, sy, sz = semiaxes
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d *= step
    out = np.empty((n, 3))
    if start is None:
        while True:
            p = (rng.random(3) * 2 - 1) * np.asarray(semiaxes) * 0.8
            if (p[0] / sx) ** 2 + (p[1] / sy) ** 2 + (p[2] / sz) ** 2 < 0.64:
                break
    else:
        p = np.asarray(start, dtype=np.float64)
    x, y, z = float(p[0]), float(p[1]), float(p[2])
    lim = 0.95 ** 2
    for i in range(n):
        dx, dy, dz = d[i]
        nx, ny, nz = x + dx, y + dy, z + dz
        if (nx / sx) ** 2 + (ny / sy) ** 2 + (nz / sz) ** 2 > lim:
            nx, ny, nz = x - dx, y - dy, z - dz      # bounce back
            if (nx / sx) ** 2 + (ny / sy) ** 2 + (nz / sz) ** 2 > lim:
                nx, ny, nz = x * 0.98, y * 0.98, z * 0.98
        x, y, z = nx, ny, nz
        out[i] = (x, y, z)
    return out


def globule_positions(design: InterphaseDesign, seed=1, radius=1.5, step=0.05):
    """Telophase-like compact start structure: per-chain confined random walk
    inside the packing radius (telophase_packing_radius = 1.5); nucleolar beads
    sit next to their NOR bead.  Quantised like a stored frame."""
    rng = np.random.default_rng(seed)
    x = np.zeros((design.n, 3))
    for s, e in design.chains:
        x[s:e] = confined_walk(e - s, rng, step, (radius, radius, radius))
    for nor, nuc in design.nucleolar_bonds:
        x[nuc] = x[nor] + rng.normal(size=3) * step
    return quantize_f32(x).astype(np.float64)


def make_polymer(n, seed=4, semiaxes=(5.0, 4.0, 3.0), bond=0.3, block=20):
    """Configs 4/5: one chain of n beads, alternating A/B blocks, confined
    random walk with the bond length as step inside the ellipsoid."""
    rng = np.random.default_rng(seed)
    a = ((np.arange(n) // block) % 2 == 0).astype(np.float64)
    b = 1.0 - a
    design = InterphaseDesign(n=n, chains=np.asarray([(0, n)], dtype=np.int32), names=["poly"],
                              a=a, b=b, mobility=np.ones(n),
                              nucleolar_bonds=np.zeros((0, 2), dtype=np.int32),
                              centromeres=np.asarray([(0, 1)], dtype=np.int32), seed=seed)
    x = confined_walk(n, rng, bond, semiaxes)
    return design, quantize_f32(x).astype(np.float64)


def make_mitotic_design(design: InterphaseDesign, cfg: MitoticConfig | None = None,
                        sisters=False):
    """Coarse-grained chains (prepare.cpp:279-313): length // coarse_graining
    beads per chain, kinetochore at the centromere midpoint; optional sister
    copies (prepare.cpp:318-363) with poles at -/+ spindle_axis (365-369)."""
    cfg = cfg or MitoticConfig()
    cg = cfg.coarse_graining
    chains, kin = [], []
    pos = 0
    for (s, e), (cs, ce) in zip(design.chains, design.centromeres):
        length = max((e - s) // cg, 3)
        mid = (int(cs) + int(ce)) // 2
        k = min((mid - int(s)) // cg, length - 1)
        chains.append((pos, pos + length))
        kin.append(pos + k)
        pos += length
    chains = np.asarray(chains, dtype=np.int32)
    kin = np.asarray(kin, dtype=np.int32)
    sis = np.zeros((0, 2), dtype=np.int32)
    if sisters:
        c = len(chains)
        chains = np.concatenate([chains, chains + pos])
        kin = np.concatenate([kin, kin + pos])
        sis = np.stack([np.arange(c), np.arange(c) + c], axis=1).astype(np.int32)
        pos *= 2
    axis = np.asarray(cfg.spindle_axis)
    return MitoticDesign(n=pos, chains=chains, kinetochores=kin, sisters=sis,
                         poles=np.stack([-axis, axis]), seed=design.seed)


def rod_positions(md: MitoticDesign, cfg: MitoticConfig | None = None, seed=1):
    """Anaphase start: randomly directed straight rods around -spindle_axis
    (SIM/stage_anatelophase/simulation_driver.cpp:221-236; numpy RNG because
    libstdc++'s normal_distribution is not portable)."""
    cfg = cfg or MitoticConfig()
    rng = np.random.default_rng(seed)
    x = np.zeros((md.n, 3))
    center0 = -np.asarray(cfg.spindle_axis)
    for s, e in md.chains:
        centroid = center0 + cfg.anaphase_start_stddev * rng.normal(size=3)
        d = rng.normal(size=3)
        stepv = cfg.bond_length * d / np.linalg.norm(d)
        p = centroid - stepv * (e - s) / 2
        for i in range(s, e):
            x[i] = p
            p = p + stepv
    return quantize_f32(x).astype(np.float64)
