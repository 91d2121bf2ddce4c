"""ctypes binding of oracle/gc_oracle.h.  TEST INFRASTRUCTURE ONLY.

May be imported only from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  `OracleSystem` mirrors the method names of
the product's `api.System` so that the same model-building code can drive both.

kind = "reference": oracle/_ref/libgc_ref.so (micromd compiled unmodified)
kind = "strict"   : oracle/_ref/libgc_ref_strict.so (same, no unsafe-math)
kind = "port"     : oracle/libgc_port.so (FP64 restatement)
"""
from __future__ import annotations

import ctypes as C
import importlib.util
import subprocess
import sys
from pathlib import Path

import numpy as np

ORACLE_DIR = Path(__file__).resolve().parent
ROOT = ORACLE_DIR.parent
LIBS = {
    "reference": ORACLE_DIR / "_ref" / "libgc_ref.so",
    "strict": ORACLE_DIR / "_ref" / "libgc_ref_strict.so",
    "port": ORACLE_DIR / "libgc_port.so",
    # the reference's literal -march=native build (timed CPU baseline only; see Makefile)
    "native": ORACLE_DIR / "_ref" / "libgc_ref_native.so",
}


def _api():
    """The product's ctypes struct definitions (shared parameter structs)."""
    name = "gc_b200_api_structs"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, ROOT / "3d-genome-cycle_b200" / "api.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


class OracleRunConfig(C.Structure):
    _fields_ = [("steps", C.c_int64), ("temperature", C.c_double), ("timestep", C.c_double),
                ("spacestep", C.c_double), ("seed", C.c_uint64), ("noise", C.POINTER(C.c_double)),
                ("interphase_callback", C.c_int32), ("energy_interval", C.c_int32),
                ("traj_out", C.POINTER(C.c_double)), ("ctx_out", C.POINTER(C.c_double)),
                ("force_out", C.POINTER(C.c_double)), ("step_offset", C.c_int64),
                ("stepper", C.c_int32), ("reserved", C.c_int32), ("last_energy", C.c_double)]


def build(kinds=("port", "reference", "strict"), quiet=True):
    """Runs oracle/Makefile for the requested libraries (reference ones only if
    /root/reference is present)."""
    targets = []
    for k in kinds:
        if k == "port":
            targets.append("port")
        elif Path("/root/reference").exists():
            targets.append({"reference": "ref", "strict": "ref_strict", "native": "ref_native"}[k])
    if targets:
        subprocess.run(["make", "-C", str(ORACLE_DIR)] + targets, check=True,
                       stdout=subprocess.DEVNULL if quiet else None)


_cache = {}


def available(kind):
    return LIBS[kind].exists()


def native_usable():
    """True when the -march=native build exists and this host has every ISA flag of the
    host it was built on (so it cannot fault on an unsupported instruction)."""
    lib = LIBS["native"]
    flags = lib.with_suffix(".flags")
    if not (lib.exists() and flags.exists()):
        return False
    need = set(flags.read_text().split())
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                return need <= set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return False


def native_arch():
    f = LIBS["native"].with_suffix(".arch")
    return f.read_text().strip() if f.exists() else "unknown"


def load(kind="reference"):
    if kind in _cache:
        return _cache[kind]
    path = LIBS[kind]
    if not path.exists():
        raise RuntimeError(f"oracle library {path} is missing (run oracle_py.build())")
    A = _api()
    lib = C.CDLL(str(path))
    P, dp, ip = C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int32)
    up = C.POINTER(C.c_uint32)
    sig = {
        "gco_kind": (C.c_char_p, []),
        "gco_create": (P, [C.c_int32]),
        "gco_destroy": (None, [P]),
        "gco_set_bead_attrs": (None, [P, dp, dp, dp]),
        "gco_set_positions": (None, [P, dp]),
        "gco_get_positions": (None, [P, dp]),
        "gco_add_pair_softcore_ab": (C.c_int, [P, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int32]),
        "gco_add_pair_softcore": (C.c_int, [P, C.c_double, C.c_double]),
        "gco_add_pair_softwell": (C.c_int, [P, C.c_double, C.c_double, C.c_double, C.c_int32, ip]),
        "gco_add_bonds": (C.c_int, [P, C.POINTER(A.BondParams), C.c_int32, ip]),
        "gco_add_triples": (C.c_int, [P, C.c_double, C.c_int32, ip]),
        "gco_set_wall": (C.c_int, [P, C.POINTER(A.WallParams)]),
        "gco_add_point_source": (C.c_int, [P, C.c_int32, dp, C.c_double, C.c_double]),
        "gco_add_kfiber": (C.c_int, [P, dp, C.c_int32, ip, dp, dp]),
        "gco_set_term_enabled": (None, [P, C.c_int32, C.c_int32]),
        "gco_set_schedule": (None, [P, C.POINTER(A.Schedule)]),
        "gco_set_context": (None, [P, C.POINTER(A.Context)]),
        "gco_get_context": (None, [P, C.POINTER(A.Context)]),
        "gco_run": (None, [P, C.POINTER(OracleRunConfig)]),
        "gco_compute_force": (None, [P, C.c_uint32, dp]),
        "gco_compute_energy": (C.c_double, [P, C.c_uint32]),
        "gco_neighbor_pairs": (C.c_size_t, [P, C.c_double, up, C.c_size_t]),
        "gco_verlet_pairs": (C.c_size_t, [P, C.c_int32, up, C.c_size_t]),
        "gco_set_contact_map": (None, [P, C.c_double, C.c_int64, C.c_int32]),
        "gco_contacts_update": (None, [P]),
        "gco_contacts_fetch": (C.c_size_t, [P, up, C.c_size_t]),
        "gco_contacts_clear": (None, [P]),
        "gco_reference_normals": (C.c_size_t, [C.c_uint64, C.c_size_t, dp]),
        "gco_potential_probe": (C.c_double, [C.c_int32, C.c_double, C.c_double, C.c_double, dp, dp]),
        "gco_bending_probe": (C.c_double, [C.c_double, dp, dp, dp]),
        "gco_ellipsoid_probe": (None, [dp, dp, dp]),
        "gco_resample_chain": (None, [C.c_int32, dp, C.c_int32, dp]),
        "gco_coarse_grain": (None, [dp, C.c_int64, C.c_int64, C.c_int64, C.c_int64, dp, dp, dp]),
        "gco_quantize_f32": (None, [C.c_size_t, dp, C.c_int32, C.POINTER(C.c_float)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _cache[kind] = lib
    return lib


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _ptr(a, typ):
    return a.ctypes.data_as(C.POINTER(typ))


class OracleSystem:
    """Single-replica FP64 system evaluated by a CPU oracle library."""

    def __init__(self, n_beads, kind="reference"):
        self.lib = load(kind)
        self.kind = kind
        self.A = _api()
        self.n = int(n_beads)
        self.n_rep = 1
        self._h = self.lib.gco_create(self.n)

    def close(self):
        if getattr(self, "_h", None):
            self.lib.gco_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_bead_attrs(self, mobility, a=None, b=None):
        m = _d(mobility)
        aa = _d(a) if a is not None else None
        bb = _d(b) if b is not None else None
        self.lib.gco_set_bead_attrs(self._h, _ptr(m, C.c_double),
                                    _ptr(aa, C.c_double) if aa is not None else None,
                                    _ptr(bb, C.c_double) if bb is not None else None)

    def set_positions(self, xyz, replica=-1):
        x = _d(xyz).reshape(self.n, 3)
        self.lib.gco_set_positions(self._h, _ptr(x, C.c_double))

    def get_positions(self, replica=0):
        x = np.empty((self.n, 3))
        self.lib.gco_get_positions(self._h, _ptr(x, C.c_double))
        return x

    def set_seeds(self, seeds):
        self.seed = int(np.asarray(seeds).ravel()[0])

    def add_pair_softcore_ab(self, eps_a, sigma_a, eps_b, sigma_b, scale_by_core=True):
        return self.lib.gco_add_pair_softcore_ab(self._h, eps_a, sigma_a, eps_b, sigma_b,
                                                 int(scale_by_core))

    def add_pair_softcore(self, eps, sigma):
        return self.lib.gco_add_pair_softcore(self._h, eps, sigma)

    def add_pair_softwell(self, eps, decay, cutoff, targets):
        t = _i(targets)
        return self.lib.gco_add_pair_softwell(self._h, eps, decay, cutoff, t.size, _ptr(t, C.c_int32))

    def add_bonds(self, pairs, kind=0, K_a=1.0, K_b=0.0, l_a=0.0, l_b=0.0, mix_ab=False, scale=False):
        p = _i(pairs).reshape(-1, 2)
        bp = self.A.BondParams(kind, int(mix_ab), int(scale), 0, K_a, K_b, l_a, l_b)
        return self.lib.gco_add_bonds(self._h, C.byref(bp), p.shape[0], _ptr(p, C.c_int32))

    def add_triples(self, triples, bending_energy):
        t = _i(triples).reshape(-1, 3)
        return self.lib.gco_add_triples(self._h, bending_energy, t.shape[0], _ptr(t, C.c_int32))

    def set_wall(self, semiaxes, eps_a, sigma_a, eps_b, sigma_b, wall_a, wall_b, outward_spring,
                 scale_by_core=True, ode_enabled=False, ode_mobility=0.0, ode_spring=(0, 0, 0)):
        w = self.A.WallParams((C.c_double * 3)(*semiaxes), eps_a, sigma_a, eps_b, sigma_b, wall_a,
                              wall_b, outward_spring, int(scale_by_core), int(ode_enabled),
                              ode_mobility, (C.c_double * 3)(*ode_spring))
        return self.lib.gco_set_wall(self._h, C.byref(w))

    def add_point_source(self, kind, point, p0, p1):
        pt = _d(point)
        return self.lib.gco_add_point_source(self._h, kind, _ptr(pt, C.c_double), p0, p1)

    def add_kfiber(self, pole, beads, spring, length):
        pl, bd, sp, ln = _d(pole), _i(beads), _d(spring), _d(length)
        return self.lib.gco_add_kfiber(self._h, _ptr(pl, C.c_double), bd.size, _ptr(bd, C.c_int32),
                                       _ptr(sp, C.c_double), _ptr(ln, C.c_double))

    def set_term_enabled(self, term, enabled):
        self.lib.gco_set_term_enabled(self._h, term, int(enabled))

    def set_schedule(self, core_init, core_tau, bond_init, bond_tau, enabled=True):
        s = self.A.Schedule(int(enabled), 0, core_init, core_tau, bond_init, bond_tau)
        self.lib.gco_set_schedule(self._h, C.byref(s))

    def get_context(self, replica=0):
        c = self.A.Context()
        self.lib.gco_get_context(self._h, C.byref(c))
        return c

    def set_context(self, replica=-1, **kw):
        c = self.get_context()
        for k, v in kw.items():
            if k == "semiaxes":
                c.semiaxes = (C.c_double * 3)(*v)
            else:
                setattr(c, k, v)
        self.lib.gco_set_context(self._h, C.byref(c))

    def set_skin(self, frac):
        pass

    def run(self, steps, timestep, temperature=1.0, spacestep=0.0, restart=True, noise=None,
            seed=None, interphase_callback=False, energy_interval=0, want_traj=False,
            want_ctx=False, want_force=False, step_offset=0, stepper=None):
        """Returns dict with optional 'traj' [steps,n,3], 'ctx' [steps,8], 'force'."""
        cfg = OracleRunConfig()
        cfg.steps = int(steps)
        cfg.temperature = temperature
        cfg.timestep = timestep
        cfg.spacestep = spacestep
        cfg.seed = int(seed if seed is not None else getattr(self, "seed", 0))
        cfg.interphase_callback = int(interphase_callback)
        cfg.energy_interval = int(energy_interval)
        cfg.step_offset = int(step_offset)
        keep = []
        out = {}
        if noise is not None:
            z = _d(noise)
            assert z.size == steps * self.n * 3
            keep.append(z)
            cfg.noise = _ptr(z, C.c_double)
        if want_traj:
            out["traj"] = np.zeros((steps, self.n, 3))
            cfg.traj_out = _ptr(out["traj"], C.c_double)
        if want_ctx:
            out["ctx"] = np.zeros((steps, 8))
            cfg.ctx_out = _ptr(out["ctx"], C.c_double)
        if want_force:
            out["force"] = np.zeros((steps, self.n, 3))
            cfg.force_out = _ptr(out["force"], C.c_double)
        if stepper is None:
            stepper = 1 if (noise is not None or want_force) else 0
        cfg.stepper = int(stepper)
        self.lib.gco_run(self._h, C.byref(cfg))
        out["last_energy"] = cfg.last_energy
        return out

    def compute_force(self, mask=0xFFFFFFFF, replica=0):
        f = np.empty((self.n, 3))
        self.lib.gco_compute_force(self._h, mask, _ptr(f, C.c_double))
        return f

    def compute_energy(self, mask=0xFFFFFFFF, replica=0):
        return self.lib.gco_compute_energy(self._h, mask)

    def neighbor_pairs(self, radius, replica=0):
        cap = max(64 * self.n, 1024)
        while True:
            out = np.empty((cap, 2), dtype=np.uint32)
            m = self.lib.gco_neighbor_pairs(self._h, radius, _ptr(out, C.c_uint32), cap)
            if m <= cap:
                return out[:m].copy()
            cap = m

    def verlet_pairs(self, term):
        cap = max(256 * self.n, 1024)
        while True:
            out = np.empty((cap, 2), dtype=np.uint32)
            m = self.lib.gco_verlet_pairs(self._h, term, _ptr(out, C.c_uint32), cap)
            if m <= cap:
                return out[:m].copy()
            cap = m

    def set_contact_map(self, distance, interval=0, scale_by_core=False, table_capacity=0):
        self.lib.gco_set_contact_map(self._h, distance, interval, int(scale_by_core))

    def contacts_update(self):
        self.lib.gco_contacts_update(self._h)

    def contacts_fetch(self, replica=0):
        m = self.lib.gco_contacts_fetch(self._h, None, 0)
        out = np.empty((m, 3), dtype=np.uint32)
        if m:
            self.lib.gco_contacts_fetch(self._h, _ptr(out, C.c_uint32), m)
        return out

    def contacts_clear(self):
        self.lib.gco_contacts_clear(self._h)


def reference_normals(seed, n, kind="reference"):
    out = np.empty(n)
    m = load(kind).gco_reference_normals(seed, n, _ptr(out, C.c_double))
    return out[:m]


def potential_probe(pot, p0, p1, p2, r, kind="reference"):
    rr = _d(r)
    f = np.zeros(3)
    e = load(kind).gco_potential_probe(pot, p0, p1, p2, _ptr(rr, C.c_double), _ptr(f, C.c_double))
    return e, f


def bending_probe(e0, rij, rjk, kind="reference"):
    a, b = _d(rij), _d(rjk)
    f = np.zeros(9)
    e = load(kind).gco_bending_probe(e0, _ptr(a, C.c_double), _ptr(b, C.c_double),
                                     _ptr(f, C.c_double))
    return e, f.reshape(3, 3)


def ellipsoid_probe(semiaxes, pt, kind="reference"):
    a, b = _d(semiaxes), _d(pt)
    out = np.zeros(8)
    load(kind).gco_ellipsoid_probe(_ptr(a, C.c_double), _ptr(b, C.c_double), _ptr(out, C.c_double))
    return dict(undefined=bool(out[0]), delta=out[1:4].copy(), strain=out[4:7].copy(),
                implicit=out[7])


def resample_chain(chain, new_n, kind="reference"):
    """resample_chain of transition_interphase.cpp:13-41 (the reference's cxx-spline)."""
    c = _d(chain).reshape(-1, 3)
    out = np.empty((int(new_n), 3))
    load(kind).gco_resample_chain(c.shape[0], _ptr(c, C.c_double), int(new_n), _ptr(out, C.c_double))
    return out


def coarse_grain(positions, source_start, source_len, cg, coarse_len, disp, kind="reference"):
    """One chromosome of transition_prometaphase.cpp:81-98 -> (target, sister)."""
    x, d = _d(positions), _d(disp)
    t, s = np.empty((int(coarse_len), 3)), np.empty((int(coarse_len), 3))
    load(kind).gco_coarse_grain(_ptr(x, C.c_double), int(source_start), int(source_len), int(cg),
                                int(coarse_len), _ptr(d, C.c_double), _ptr(t, C.c_double),
                                _ptr(s, C.c_double))
    return t, s


def quantize_f32(x, bits=16, kind="reference"):
    """simulation_store.cpp:22-32 + the float store of save_positions."""
    v = _d(x)
    out = np.empty(v.shape, dtype=np.float32)
    load(kind).gco_quantize_f32(v.size, _ptr(v, C.c_double), int(bits), _ptr(out, C.c_float))
    return out
