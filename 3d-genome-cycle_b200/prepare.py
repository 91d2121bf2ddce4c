"""Topology preparation as a library: chains.tsv -> bead tables,
nucleolar beads, coarse-grained mitotic chains, kinetochore indices, sister pairs, poles and
the per-stage seeds -- what the reference's `prepare init` writes into the trajectory file.

Restated from (SIM = src/simulation):
  chains.tsv reader             SIM/stage_prepare/chains.cpp:14-63 (header row, columns
                                chain start end a b tags; contiguous rows of one name = a chain)
  define_chains                 SIM/stage_prepare/prepare.cpp:168-218 (tag priority anor, bnor,
                                cen, A, B, u; active NORs remembered)
  define_nucleolar_particles    prepare.cpp:221-238
  define_anatelophase_chains    prepare.cpp:241-313 (length / coarse_graining beads per chain,
                                kinetochore bin = (centromere midpoint - start) / coarse_graining)
  define_prometaphase_chains    prepare.cpp:318-369 (target / sister chains INTERLEAVED:
                                target c at 2 start_c, its "-copy" right behind it; poles at
                                -/+ spindle_axis)
  write_seeds                   prepare.cpp:549-562: std::seed_seq{master}.generate(3 words)
                                = seeds of anaphase, interphase, prometaphase
and the drivers' use of those seeds: `std::mt19937_64 _random(design.seed)` whose successive
outputs seed every md::simulate_brownian_dynamics call
(SIM/stage_interphase/simulation_driver.cpp:19, ..._relaxation.cpp:53, ..._interphase.cpp:61;
SIM/stage_anatelophase/simulation_driver.cpp:40,266,307; SIM/stage_prometaphase/...:18,234).
std::seed_seq and std::mt19937_64 are fully specified by the C++ standard; they are restated
here so that `run_seeds(master, ...)` hands gc_set_seeds the same 64-bit values a reference
`scripts/simulate -s master` run passes to its integrator (tests/test_prepare.py checks them
against libstdc++ itself).  The noise STREAM still differs: Philox here, sfc64 + ziggurat there.
"""
from __future__ import annotations

import io
from dataclasses import dataclass, field

import numpy as np

from .synth import InterphaseConfig, InterphaseDesign, MitoticConfig, MitoticDesign

TAG_TYPES = ("anor", "bnor", "cen", "A", "B", "u")        # priority order of define_chains
NUCLEOLUS, UNKNOWN = "nucleolus", "unknown"


# ---------------------------------------------------------------------------
# chains.tsv
# ---------------------------------------------------------------------------
@dataclass
class Bead:
    bin_start: int
    bin_end: int
    a_factor: float
    b_factor: float
    tags: str


@dataclass
class Chain:
    name: str
    beads: list = field(default_factory=list)


COLUMNS = ("chain", "start", "end", "a", "b", "tags")


def parse_chains_tsv(text: str) -> list:
    """load_chains (chains.cpp:14-63): tab-separated, one header row, six columns;
    contiguous records with the same chain name form one chain."""
    lines = [ln for ln in text.splitlines() if ln.strip() != ""]
    if not lines:
        raise ValueError("cannot load chain definitions: header is expected but not seen")
    header = lines[0].split("\t")
    if tuple(header) != COLUMNS:
        raise ValueError(f"cannot load chain definitions: header {header} != {list(COLUMNS)}")
    chains: list = []
    current = Chain(name="")
    for ln, line in enumerate(lines[1:], start=2):
        f = line.split("\t")
        if len(f) != 6:
            raise ValueError(f"cannot load chain definitions: line {ln}: expected 6 fields, got {len(f)}")
        try:
            start, end = int(f[1]), int(f[2])
            a, b = float(f[3]), float(f[4])
        except ValueError as ex:
            raise ValueError(f"cannot load chain definitions: line {ln}: {ex}") from None
        if start < 0 or end < 0 or start >= 2 ** 32 or end >= 2 ** 32:
            raise ValueError(f"cannot load chain definitions: line {ln}: bin outside uint32")
        if f[0] != current.name:
            if current.beads:
                chains.append(current)
            current = Chain(name=f[0])
        current.beads.append(Bead(start, end, a, b, f[5]))
    if current.beads:
        chains.append(current)
    return chains


def load_chains_tsv(path) -> list:
    with open(path, "r", encoding="utf-8") as fh:
        return parse_chains_tsv(fh.read())


def format_chains_tsv(chains) -> str:
    out = io.StringIO()
    out.write("\t".join(COLUMNS) + "\n")
    for ch in chains:
        for bd in ch.beads:
            out.write(f"{ch.name}\t{bd.bin_start}\t{bd.bin_end}\t{bd.a_factor!r}\t{bd.b_factor!r}\t{bd.tags}\n")
    return out.getvalue()


def check_tag(tags: str, tag: str) -> bool:
    """prepare.cpp:143-165: comma-delimited exact field match."""
    return tag in tags.split(",")


def chains_from_design(design: InterphaseDesign, bead_bp=100000) -> list:
    """The chains.tsv rows a synthetic design corresponds to (synth.make_diploid_design):
    tags from the (a, b) factors as src/annotate.py:138-142 writes them, `cen` on the
    centromeric block, `anor` on the NOR beads."""
    nors = set(int(i) for i in np.unique(design.nucleolar_bonds[:, 0])) if len(design.nucleolar_bonds) else set()
    chains = []
    for name, (s, e), (cs, ce) in zip(design.names, design.chains, design.centromeres):
        ch = Chain(name=name)
        for i in range(int(s), int(e)):
            a, b = float(design.a[i]), float(design.b[i])
            tag = "A" if a > b else ("B" if b > a else "u")
            if i in nors:
                tag = "anor," + tag
            elif cs <= i < ce:
                tag = "cen," + tag
            k = i - int(s)
            ch.beads.append(Bead(k * bead_bp, (k + 1) * bead_bp, a, b, tag))
        chains.append(ch)
    return chains


# ---------------------------------------------------------------------------
# std::seed_seq / std::mt19937_64 (ISO C++ [rand.util.seedseq], [rand.eng.mers])
# ---------------------------------------------------------------------------
def seed_seq_generate(seeds, n: int) -> list:
    """std::seed_seq{seeds...}.generate(begin, begin + n) -> n 32-bit words."""
    M32 = 0xFFFFFFFF
    v = [int(s) & M32 for s in seeds]
    if n == 0:
        return []
    s = len(v)
    out = [0x8B8B8B8B] * n
    t = 11 if n >= 623 else 7 if n >= 68 else 5 if n >= 39 else 3 if n >= 7 else (n - 1) // 2
    p = (n - t) // 2
    q = p + t
    m = max(s + 1, n)

    def T(x):
        return x ^ (x >> 27)

    for k in range(m):
        r1 = (1664525 * T(out[k % n] ^ out[(k + p) % n] ^ out[(k - 1) % n])) & M32
        if k == 0:
            r2 = r1 + s
        elif k <= s:
            r2 = r1 + (k % n) + v[k - 1]
        else:
            r2 = r1 + (k % n)
        r2 &= M32
        out[(k + p) % n] = (out[(k + p) % n] + r1) & M32
        out[(k + q) % n] = (out[(k + q) % n] + r2) & M32
        out[k % n] = r2
    for k in range(m, m + n):
        r3 = (1566083941 * T((out[k % n] + out[(k + p) % n] + out[(k - 1) % n]) & M32)) & M32
        r4 = (r3 - (k % n)) & M32
        out[(k + p) % n] ^= r3
        out[(k + q) % n] ^= r4
        out[k % n] = r4
    return out


class MT19937_64:
    """std::mt19937_64."""
    NN, MM = 312, 156
    MATRIX_A, UM, LM = 0xB5026F5AA96619E9, 0xFFFFFFFF80000000, 0x7FFFFFFF
    M64 = 0xFFFFFFFFFFFFFFFF

    def __init__(self, seed: int = 5489):
        self.mt = [0] * self.NN
        self.mt[0] = int(seed) & self.M64
        for i in range(1, self.NN):
            x = self.mt[i - 1]
            self.mt[i] = (6364136223846793005 * (x ^ (x >> 62)) + i) & self.M64
        self.idx = self.NN

    def __call__(self) -> int:
        if self.idx >= self.NN:
            mt, NN, MM = self.mt, self.NN, self.MM
            for i in range(NN):
                x = (mt[i] & self.UM) | (mt[(i + 1) % NN] & self.LM)
                mt[i] = mt[(i + MM) % NN] ^ (x >> 1) ^ (self.MATRIX_A if x & 1 else 0)
            self.idx = 0
        x = self.mt[self.idx]
        self.idx += 1
        x ^= (x >> 29) & 0x5555555555555555
        x ^= (x << 17) & 0x71D67FFFEDA60000
        x ^= (x << 37) & 0xFFF7EEE000000000
        x ^= x >> 43
        return x & self.M64


def stage_seeds(master_seed: int) -> dict:
    """write_seeds (prepare.cpp:549-562)."""
    a, i, p = seed_seq_generate([master_seed], 3)
    return {"anaphase": a, "interphase": i, "prometaphase": p}


def run_seeds(master_seed: int) -> dict:
    """The 64-bit seed of every md::simulate_brownian_dynamics call of a reference run started
    with `-s master_seed`, in call order.  The anaphase driver first consumes its generator for
    the random start rods when no initial structure exists (std::normal_distribution on the
    same mt19937_64, not portable: simulation_driver.cpp:221-236); the values here are those
    of a run that STARTS from a stored structure (the daughter cell of transition cycle), and
    of relaxation / interphase / prometaphase, which never draw anything else."""
    s = stage_seeds(master_seed)
    ra, ri, rp = MT19937_64(s["anaphase"]), MT19937_64(s["interphase"]), MT19937_64(s["prometaphase"])
    return {"anaphase": ra(), "telophase": ra(), "relaxation": ri(), "interphase": ri(),
            "prometaphase": rp()}


# ---------------------------------------------------------------------------
# init_pipeline
# ---------------------------------------------------------------------------
@dataclass
class Prepared:
    interphase: InterphaseDesign
    anatelophase: MitoticDesign
    prometaphase: MitoticDesign
    particle_types: list            # interphase, per bead
    mitotic_types: list             # anatelophase, per bead: "arm" / "kinetochore"
    nor_indices: np.ndarray
    seeds: dict
    chain_names: list
    prometaphase_chain_names: list


def prepare(chains, icfg: InterphaseConfig | None = None, mcfg: MitoticConfig | None = None,
            master_seed: int = 0) -> Prepared:
    """init_pipeline::run (prepare.cpp:128-141) without the file writes."""
    icfg = icfg or InterphaseConfig()
    mcfg = mcfg or MitoticConfig()
    # define_chains
    a, b, types, nors, ranges, names = [], [], [], [], [], []
    for ch in chains:
        start = len(a)
        for bd in ch.beads:
            t = UNKNOWN
            for tag in TAG_TYPES:
                if check_tag(bd.tags, tag):
                    t = tag
                    break
            if t == "anor":
                nors.append(len(a))
            a.append(bd.a_factor)
            b.append(bd.b_factor)
            types.append(t)
        ranges.append((start, len(a)))
        names.append(ch.name)
    # define_nucleolar_particles
    bonds = []
    for nor in nors:
        for _ in range(icfg.nucleolus_bead_count):
            bonds.append((nor, len(a)))
            a.append(icfg.nucleolus_ab_factor[0])
            b.append(icfg.nucleolus_ab_factor[1])
            types.append(NUCLEOLUS)
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    n = len(a)
    # centromeric range per chain (define_anatelophase_chains, first part)
    cens = []
    for s, e in ranges:
        cs, ce, seen = s, e, False
        for i in range(s, e):
            if types[i] == "cen":
                if not seen:
                    cs, seen = i, True
                ce = i + 1
        cens.append((cs, ce))
    # mobilities (simulation_driver_particles.cpp:19-34)
    mob = np.zeros(n)
    for s, e in ranges:
        mob[s:e] = np.where(a[s:e] >= b[s:e], icfg.a_core_mobility, icfg.b_core_mobility)
    bonds_arr = np.asarray(bonds, dtype=np.int32).reshape(-1, 2)
    if len(bonds_arr):
        mob[bonds_arr[:, 1]] = icfg.nucleolus_mobility
    seeds = stage_seeds(master_seed)
    inter = InterphaseDesign(n=n, chains=np.asarray(ranges, dtype=np.int32).reshape(-1, 2), names=names, a=a,
                             b=b, mobility=mob, nucleolar_bonds=bonds_arr,
                             centromeres=np.asarray(cens, dtype=np.int32).reshape(-1, 2),
                             seed=seeds["interphase"])
    # define_anatelophase_chains
    cg = int(mcfg.coarse_graining)
    mchains, kin, mtypes = [], [], []
    for (s, e), (cs, ce) in zip(ranges, cens):
        coarse_len = (e - s) // cg
        start = len(mtypes)
        koff = ((cs + ce) // 2 - s) // cg
        k_index = None
        for bin_ in range(coarse_len):
            if bin_ == koff:
                k_index = len(mtypes)
                mtypes.append("kinetochore")
            else:
                mtypes.append("arm")
        if k_index is None:
            raise ValueError(f"chain [{s}, {e}): kinetochore bin {koff} outside its {coarse_len} coarse beads "
                             "(the reference dereferences an empty optional here)")
        mchains.append((start, start + coarse_len))
        kin.append(k_index)
    axis = np.asarray(mcfg.spindle_axis, dtype=np.float64)
    ana = MitoticDesign(n=len(mtypes), chains=np.asarray(mchains, dtype=np.int32).reshape(-1, 2),
                        kinetochores=np.asarray(kin, dtype=np.int32), poles=np.stack([-axis, axis]),
                        seed=seeds["anaphase"])
    # define_prometaphase_chains: target c at 2 * start_c, its copy right behind it
    pchains, pkin, sisters, pnames = [], [], [], []
    for c, ((s, e), k) in enumerate(zip(mchains, kin)):
        length, koff = e - s, k - s
        ts = 2 * s
        pchains.append((ts, ts + length))
        pchains.append((ts + length, ts + 2 * length))
        pkin.extend([ts + koff, ts + length + koff])
        sisters.append((2 * c, 2 * c + 1))
        pnames.extend([names[c], names[c] + "-copy"])
    pro = MitoticDesign(n=2 * len(mtypes), chains=np.asarray(pchains, dtype=np.int32).reshape(-1, 2),
                        kinetochores=np.asarray(pkin, dtype=np.int32),
                        sisters=np.asarray(sisters, dtype=np.int32).reshape(-1, 2),
                        poles=np.stack([-axis, axis]), seed=seeds["prometaphase"])
    return Prepared(interphase=inter, anatelophase=ana, prometaphase=pro, particle_types=types,
                    mitotic_types=mtypes, nor_indices=np.asarray(nors, dtype=np.int32), seeds=seeds,
                    chain_names=names, prometaphase_chain_names=pnames)
