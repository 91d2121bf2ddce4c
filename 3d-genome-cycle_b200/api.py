"""ctypes binding of the C-ABI in include/gc_b200.h.

The product path is libgc_b200.so (CUDA, sm_100a).  There is no CPU fallback:
if the library is missing or no GPU is present this module raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

PKG_DIR = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("GC_B200_LIB", PKG_DIR / "libgc_b200.so"))

GC_BOND_SEMISPRING, GC_BOND_HARMONIC, GC_BOND_SPRING = 0, 1, 2
GC_SOURCE_SEMISPRING, GC_SOURCE_FORCE_FLUX = 0, 1
GC_MASK_ALL = 0xFFFFFFFF


class BondParams(C.Structure):
    _fields_ = [("kind", C.c_int32), ("mix_ab", C.c_int32), ("scale", C.c_int32),
                ("reserved", C.c_int32), ("K_a", C.c_double), ("K_b", C.c_double),
                ("l_a", C.c_double), ("l_b", C.c_double)]


class WallParams(C.Structure):
    _fields_ = [("semiaxes", C.c_double * 3), ("eps_a", C.c_double), ("sigma_a", C.c_double),
                ("eps_b", C.c_double), ("sigma_b", C.c_double), ("wall_a", C.c_double),
                ("wall_b", C.c_double), ("outward_spring", C.c_double),
                ("scale_by_core", C.c_int32), ("ode_enabled", C.c_int32),
                ("ode_mobility", C.c_double), ("ode_spring", C.c_double * 3)]


class Schedule(C.Structure):
    _fields_ = [("enabled", C.c_int32), ("reserved", C.c_int32), ("core_init", C.c_double),
                ("core_tau", C.c_double), ("bond_init", C.c_double), ("bond_tau", C.c_double)]


class Context(C.Structure):
    _fields_ = [("time", C.c_double), ("semiaxes", C.c_double * 3), ("core_scale", C.c_double),
                ("bond_scale", C.c_double), ("contact_distance", C.c_double),
                ("axial_reaction", C.c_double * 3), ("reaction_force", C.c_double),
                ("last_timestep", C.c_double), ("step", C.c_int64), ("rebuilds", C.c_int64)]

    def as_dict(self):
        return dict(time=self.time, semiaxes=list(self.semiaxes), core_scale=self.core_scale,
                    bond_scale=self.bond_scale, contact_distance=self.contact_distance,
                    axial_reaction=list(self.axial_reaction), reaction_force=self.reaction_force,
                    last_timestep=self.last_timestep, step=self.step, rebuilds=self.rebuilds)


class RunConfig(C.Structure):
    _fields_ = [("steps", C.c_int64), ("temperature", C.c_double), ("timestep", C.c_double),
                ("spacestep", C.c_double), ("restart", C.c_int32), ("reserved", C.c_int32),
                ("noise", C.POINTER(C.c_double))]


# every symbol include/gc_b200.h declares: (name, restype, argtypes)
_P = C.c_void_p
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
SYMBOLS = [
    ("gc_create", C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.POINTER(_P)]),
    ("gc_destroy", C.c_int, [_P]),
    ("gc_last_error", C.c_char_p, [_P]),
    ("gc_version", C.c_char_p, []),
    ("gc_set_bead_attrs", C.c_int, [_P, _dp, _dp, _dp]),
    ("gc_set_positions", C.c_int, [_P, C.c_int32, _dp]),
    ("gc_get_positions", C.c_int, [_P, C.c_int32, _dp]),
    ("gc_set_positions_f32", C.c_int, [_P, C.c_int32, C.POINTER(C.c_float)]),
    ("gc_get_positions_f32", C.c_int, [_P, C.c_int32, C.POINTER(C.c_float)]),
    ("gc_get_frame", C.c_int, [_P, C.c_int32, C.c_int32, C.POINTER(C.c_float)]),
    ("gc_set_positions_all_f32", C.c_int, [_P, C.POINTER(C.c_float)]),
    ("gc_get_positions_all_f32", C.c_int, [_P, C.POINTER(C.c_float)]),
    ("gc_host_alloc", C.c_int, [C.c_size_t, C.POINTER(C.c_void_p)]),
    ("gc_host_free", C.c_int, [C.c_void_p]),
    ("gc_frame_begin", C.c_int, [_P, C.c_int32, C.c_int32]),
    ("gc_frame_end", C.c_int, [_P, C.POINTER(C.c_float)]),
    ("gc_get_contexts", C.c_int, [_P, C.POINTER(Context)]),
    ("gc_set_step", C.c_int, [_P, C.c_int32, C.c_int64]),
    ("gc_compute_energy_all", C.c_int, [_P, C.c_uint32, _dp]),
    ("gc_set_seeds", C.c_int, [_P, C.POINTER(C.c_uint64)]),
    ("gc_add_pair_softcore_ab", C.c_int, [_P, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int32]),
    ("gc_add_pair_softcore", C.c_int, [_P, C.c_double, C.c_double]),
    ("gc_add_pair_softwell", C.c_int, [_P, C.c_double, C.c_double, C.c_double, C.c_int32, _ip]),
    ("gc_add_bonds", C.c_int, [_P, C.POINTER(BondParams), C.c_int32, _ip]),
    ("gc_add_triples", C.c_int, [_P, C.c_double, C.c_int32, _ip]),
    ("gc_set_wall", C.c_int, [_P, C.POINTER(WallParams)]),
    ("gc_add_point_source", C.c_int, [_P, C.c_int32, _dp, C.c_double, C.c_double]),
    ("gc_add_kfiber", C.c_int, [_P, _dp, C.c_int32, _ip, _dp, _dp]),
    ("gc_set_term_enabled", C.c_int, [_P, C.c_int32, C.c_int32]),
    ("gc_set_schedule", C.c_int, [_P, C.POINTER(Schedule)]),
    ("gc_set_context", C.c_int, [_P, C.c_int32, C.POINTER(Context)]),
    ("gc_get_context", C.c_int, [_P, C.c_int32, C.POINTER(Context)]),
    ("gc_run", C.c_int, [_P, C.POINTER(RunConfig)]),
    ("gc_set_skin", C.c_int, [_P, C.c_double]),
    ("gc_debug_noise", C.c_int, [_P, C.c_int32, C.c_int64, C.POINTER(C.c_float)]),
    ("gc_compute_force", C.c_int, [_P, C.c_int32, C.c_uint32, _dp]),
    ("gc_compute_energy", C.c_int, [_P, C.c_int32, C.c_uint32, _dp]),
    ("gc_neighbor_pairs", C.c_int, [_P, C.c_int32, C.c_double, C.POINTER(C.c_uint32), C.c_size_t,
                                    C.POINTER(C.c_size_t)]),
    ("gc_set_contact_map", C.c_int, [_P, C.c_double, C.c_int64, C.c_int32, C.c_size_t]),
    ("gc_contacts_update", C.c_int, [_P]),
    ("gc_contacts_fetch", C.c_int, [_P, C.c_int32, C.POINTER(C.c_uint32), C.c_size_t,
                                    C.POINTER(C.c_size_t)]),
    ("gc_contacts_clear", C.c_int, [_P]),
    ("gc_contacts_reduce_device", C.c_int, [_P, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                            C.POINTER(C.c_size_t)]),
    ("gc_contacts_merge_device", C.c_int, [_P, C.c_void_p, C.c_void_p, C.c_size_t,
                                           C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                           C.POINTER(C.c_size_t)]),
    ("gc_decomp_configure", C.c_int, [_P, C.c_int32, C.c_int32]),
    ("gc_decomp_set_halo", C.c_int, [_P, C.c_int32, C.c_int32]),
    ("gc_decomp_buffers", C.c_int, [_P] + [C.POINTER(C.c_void_p)] * 6 + [C.POINTER(C.c_size_t)]),
    ("gc_decomp_step_pack", C.c_int, [_P]),
    ("gc_decomp_owned", C.c_int, [_P, C.c_void_p]),
    ("gc_decomp_host_advance", C.c_int, [_P, C.c_int64]),
    ("gc_decomp_set_owned_bound", C.c_int, [_P, C.c_int32]),
    ("gc_decomp_recv_buffers", C.c_int, [_P, C.POINTER(C.c_void_p)]),
    ("gc_decomp_peer_export", C.c_int, [_P, C.c_void_p]),
    ("gc_decomp_peer_attach_ipc", C.c_int, [_P, C.c_void_p, C.c_void_p]),
    ("gc_decomp_peer_attach", C.c_int, [_P, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    ("gc_decomp_peer_detach", C.c_int, [_P]),
    ("gc_decomp_step_compute", C.c_int, [_P]),
    ("gc_decomp_step_commit", C.c_int, [_P]),
    ("gc_transition_interphase", C.c_int, [_P, C.c_int32, _P, C.c_int32, C.c_int32, _ip, _ip, C.c_int32, _ip, C.c_int32]),
    ("gc_transition_prometaphase", C.c_int, [_P, C.c_int32, _P, C.c_int32, C.c_int32, _ip, _ip, _ip, C.c_int32, _dp, C.c_int32]),
    ("gc_transition_cycle", C.c_int, [_P, C.c_int32, _P, C.c_int32, C.c_int32, _ip, _ip, _dp, C.c_int32]),
    ("gc_obs_create", C.c_int, [C.c_int32, C.c_int32, C.c_int32, _ip, C.POINTER(_P)]),
    ("gc_obs_destroy", C.c_int, [_P]),
    ("gc_obs_last_error", C.c_char_p, [_P]),
    ("gc_obs_set_bead_map", C.c_int, [_P, C.c_int32, _ip, _ip]),
    ("gc_obs_clear", C.c_int, [_P]),
    ("gc_obs_add_runs_device", C.c_int, [_P, C.c_void_p, C.c_void_p, C.c_size_t]),
    ("gc_obs_add_triples", C.c_int, [_P, C.POINTER(C.c_uint32), C.c_size_t]),
    ("gc_obs_get_matrix", C.c_int, [_P, C.POINTER(C.c_float)]),
    ("gc_obs_set_matrix", C.c_int, [_P, C.POINTER(C.c_float)]),
    ("gc_obs_ps_runs_device", C.c_int, [_P, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int32, C.POINTER(C.c_uint64)]),
    ("gc_obs_expected", C.c_int, [_P, _dp, C.c_int32, _dp]),
    ("gc_obs_max_separation", C.c_int, [_P, _ip]),
    ("gc_obs_pc1", C.c_int, [_P, C.c_int32, C.c_int32, C.c_double, C.c_int32, C.POINTER(C.c_float),
                             C.POINTER(C.c_float), _dp, _ip]),
    ("gc_obs_nci", C.c_int, [_P, C.POINTER(C.c_float)]),
    ("gc_obs_type_enrichment", C.c_int, [_P, _ip, C.c_int32, C.c_int32, _dp]),
    ("gc_radius_of_gyration", C.c_int, [_P, C.c_int32, C.c_int32, _ip, _dp]),
    ("gc_radius_of_gyration_noalloc", C.c_int, [_P, C.c_int32, C.c_int32, _ip, _dp]),
    ("gc_debug_grid", C.c_int, [_P, C.c_int32, _dp]),
    ("gc_profile_enable", C.c_int, [_P, C.c_int32]),
    ("gc_profile_read", C.c_int, [_P, C.c_int32, C.c_char_p, C.c_size_t, C.POINTER(C.c_int64), _dp]),
    ("gc_measure_fp32_peak", C.c_int, [C.c_int32, _dp]),
    ("gc_launch_count", C.c_int, [_P, C.POINTER(C.c_int64)]),
    ("gc_synchronize", C.c_int, [_P]),
    ("gc_stream", C.c_int, [_P, C.POINTER(C.c_void_p)]),
]

_lib = None


def load_library(path: os.PathLike | None = None) -> C.CDLL:
    """Loads libgc_b200.so and types every exported symbol.  Raises if absent."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = Path(path) if path else LIB_PATH
    if not p.exists():
        raise RuntimeError(f"{p} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
    lib = C.CDLL(str(p))
    for name, res, args in SYMBOLS:
        fn = getattr(lib, name)   # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib


def measure_fp32_peak(device=0):
    """Measured FP32 FFMA peak of the device in TFLOP/s."""
    v = C.c_double()
    rc = load_library().gc_measure_fp32_peak(int(device), C.byref(v))
    if rc != 0:
        raise GcError(rc, "gc_measure_fp32_peak failed")
    return v.value


class GcError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"gc_b200 error {code}: {message}")
        self.code = code


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _ptr(a, typ):
    return a.ctypes.data_as(C.POINTER(typ))


class _DeviceArray:
    """Raw device pointer -> __cuda_array_interface__ (zero-copy view for torch)."""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 3}


class System:
    """R replicas of an N-bead system on one GPU (thin wrapper over gc_ctx)."""

    def __init__(self, n_beads: int, n_replicas: int = 1, device: int = 0):
        self.lib = load_library()
        self.n = int(n_beads)
        self.n_rep = int(n_replicas)
        self._h = _P()
        rc = self.lib.gc_create(self.n, self.n_rep, int(device), C.byref(self._h))
        if rc != 0:
            msg = self.lib.gc_last_error(self._h).decode() if self._h else "no CUDA device"
            raise GcError(rc, msg)

    def close(self):
        if getattr(self, "_h", None):
            self.lib.gc_destroy(self._h)
            self._h = None
            for ptr in getattr(self, "_host_allocs", []):
                self.lib.gc_host_free(ptr)
            self._host_allocs = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc < 0:
            raise GcError(rc, self.lib.gc_last_error(self._h).decode())
        return rc

    # ---- beads -----------------------------------------------------------
    def set_bead_attrs(self, mobility, a=None, b=None):
        m = _d(mobility)
        aa = _d(a) if a is not None else None
        bb = _d(b) if b is not None else None
        self._ck(self.lib.gc_set_bead_attrs(self._h, _ptr(m, C.c_double),
                                            _ptr(aa, C.c_double) if aa is not None else None,
                                            _ptr(bb, C.c_double) if bb is not None else None))

    def set_positions(self, xyz, replica=-1):
        x = _d(xyz).reshape(self.n, 3)
        self._ck(self.lib.gc_set_positions(self._h, replica, _ptr(x, C.c_double)))

    def get_positions(self, replica=0):
        x = np.empty((self.n, 3), dtype=np.float64)
        self._ck(self.lib.gc_get_positions(self._h, replica, _ptr(x, C.c_double)))
        return x

    def set_positions_f32(self, xyz, replica=-1):
        x = np.ascontiguousarray(xyz, dtype=np.float32).reshape(self.n, 3)
        self._ck(self.lib.gc_set_positions_f32(self._h, replica, _ptr(x, C.c_float)))

    def get_positions_f32(self, replica=0):
        x = np.empty((self.n, 3), dtype=np.float32)
        self._ck(self.lib.gc_get_positions_f32(self._h, replica, _ptr(x, C.c_float)))
        return x

    def get_frame(self, replica=0, fraction_bits=16):
        """Quantised f32 frame as the reference stores it; replica = -1: [R, n, 3]."""
        nrep = self.n_rep if replica < 0 else 1
        out = np.empty((nrep, self.n, 3), dtype=np.float32)
        self._ck(self.lib.gc_get_frame(self._h, int(replica), int(fraction_bits), _ptr(out, C.c_float)))
        return out if replica < 0 else out[0]

    # ---- whole-ensemble transfers (one copy per direction) ---------------------
    def host_buffer(self, shape=None):
        """Page-locked float32 array (default [R, n, 3]) for the *_all transfers; it is
        freed with the System."""
        shape = tuple(shape) if shape is not None else (self.n_rep, self.n, 3)
        count = int(np.prod(shape))
        ptr = C.c_void_p()
        if self.lib.gc_host_alloc(count * 4, C.byref(ptr)) != 0:
            raise GcError(-2, "gc_host_alloc failed")
        self._host_allocs = getattr(self, "_host_allocs", [])
        self._host_allocs.append(ptr)
        buf = (C.c_float * count).from_address(ptr.value)
        return np.frombuffer(buf, dtype=np.float32).reshape(shape)

    def set_positions_all_f32(self, xyz):
        x = xyz if (isinstance(xyz, np.ndarray) and xyz.dtype == np.float32 and
                    xyz.flags.c_contiguous) else np.ascontiguousarray(xyz, dtype=np.float32)
        assert x.size == self.n_rep * self.n * 3
        self._ck(self.lib.gc_set_positions_all_f32(self._h, _ptr(x, C.c_float)))
        self._keep_upload = x            # a page-locked source is read asynchronously

    def get_positions_all_f32(self, out=None):
        if out is None:
            out = np.empty((self.n_rep, self.n, 3), dtype=np.float32)
        assert out.dtype == np.float32 and out.flags.c_contiguous and out.size == self.n_rep * self.n * 3
        self._ck(self.lib.gc_get_positions_all_f32(self._h, _ptr(out, C.c_float)))
        return out

    def frame_begin(self, replica=-1, fraction_bits=16):
        """Starts an asynchronous quantised frame; steps queued afterwards overlap its copy."""
        self._frame_shape = (self.n_rep if replica < 0 else 1, self.n, 3)
        self._ck(self.lib.gc_frame_begin(self._h, int(replica), int(fraction_bits)))

    def frame_end(self):
        out = np.empty(self._frame_shape, dtype=np.float32)
        self._ck(self.lib.gc_frame_end(self._h, _ptr(out, C.c_float)))
        return out

    def get_contexts(self):
        arr = (Context * self.n_rep)()
        self._ck(self.lib.gc_get_contexts(self._h, arr))
        return list(arr)

    def compute_energy_all(self, mask=GC_MASK_ALL):
        e = np.empty(self.n_rep, dtype=np.float64)
        self._ck(self.lib.gc_compute_energy_all(self._h, mask, _ptr(e, C.c_double)))
        return e

    def set_seeds(self, seeds):
        s = np.ascontiguousarray(seeds, dtype=np.uint64)
        assert s.size == self.n_rep
        self._ck(self.lib.gc_set_seeds(self._h, _ptr(s, C.c_uint64)))

    # ---- terms -----------------------------------------------------------
    def add_pair_softcore_ab(self, eps_a, sigma_a, eps_b, sigma_b, scale_by_core=True):
        return self._ck(self.lib.gc_add_pair_softcore_ab(self._h, eps_a, sigma_a, eps_b, sigma_b,
                                                         int(scale_by_core)))

    def add_pair_softcore(self, eps, sigma):
        return self._ck(self.lib.gc_add_pair_softcore(self._h, eps, sigma))

    def add_pair_softwell(self, eps, decay, cutoff, targets):
        t = _i(targets)
        return self._ck(self.lib.gc_add_pair_softwell(self._h, eps, decay, cutoff, t.size,
                                                      _ptr(t, C.c_int32)))

    def add_bonds(self, pairs, kind=GC_BOND_SEMISPRING, K_a=1.0, K_b=0.0, l_a=0.0, l_b=0.0,
                  mix_ab=False, scale=False):
        p = _i(pairs).reshape(-1, 2)
        bp = BondParams(kind, int(mix_ab), int(scale), 0, K_a, K_b, l_a, l_b)
        return self._ck(self.lib.gc_add_bonds(self._h, C.byref(bp), p.shape[0], _ptr(p, C.c_int32)))

    def add_triples(self, triples, bending_energy):
        t = _i(triples).reshape(-1, 3)
        return self._ck(self.lib.gc_add_triples(self._h, bending_energy, t.shape[0],
                                                _ptr(t, C.c_int32)))

    def set_wall(self, semiaxes, eps_a, sigma_a, eps_b, sigma_b, wall_a, wall_b, outward_spring,
                 scale_by_core=True, ode_enabled=False, ode_mobility=0.0, ode_spring=(0, 0, 0)):
        w = WallParams((C.c_double * 3)(*semiaxes), eps_a, sigma_a, eps_b, sigma_b, wall_a, wall_b,
                       outward_spring, int(scale_by_core), int(ode_enabled), ode_mobility,
                       (C.c_double * 3)(*ode_spring))
        return self._ck(self.lib.gc_set_wall(self._h, C.byref(w)))

    def add_point_source(self, kind, point, p0, p1):
        pt = _d(point)
        return self._ck(self.lib.gc_add_point_source(self._h, kind, _ptr(pt, C.c_double), p0, p1))

    def add_kfiber(self, pole, beads, spring, length):
        pl, bd, sp, ln = _d(pole), _i(beads), _d(spring), _d(length)
        return self._ck(self.lib.gc_add_kfiber(self._h, _ptr(pl, C.c_double), bd.size,
                                               _ptr(bd, C.c_int32), _ptr(sp, C.c_double),
                                               _ptr(ln, C.c_double)))

    def set_term_enabled(self, term, enabled):
        self._ck(self.lib.gc_set_term_enabled(self._h, term, int(enabled)))

    # ---- schedule / context ------------------------------------------------
    def set_schedule(self, core_init, core_tau, bond_init, bond_tau, enabled=True):
        s = Schedule(int(enabled), 0, core_init, core_tau, bond_init, bond_tau)
        self._ck(self.lib.gc_set_schedule(self._h, C.byref(s)))

    def get_context(self, replica=0) -> Context:
        c = Context()
        self._ck(self.lib.gc_get_context(self._h, replica, C.byref(c)))
        return c

    def set_context(self, replica=-1, **kw):
        c = self.get_context(max(replica, 0))
        for k, v in kw.items():
            if k == "semiaxes":
                c.semiaxes = (C.c_double * 3)(*v)
            else:
                setattr(c, k, v)
        self._ck(self.lib.gc_set_context(self._h, replica, C.byref(c)))

    def set_step(self, step, replica=-1):
        """Resume a logical run at `step` (use with run(..., restart=False))."""
        self._ck(self.lib.gc_set_step(self._h, int(replica), int(step)))

    def set_skin(self, frac):
        self._ck(self.lib.gc_set_skin(self._h, float(frac)))

    # ---- dynamics ----------------------------------------------------------
    def run(self, steps, timestep, temperature=1.0, spacestep=0.0, restart=True, noise=None):
        cfg = RunConfig(int(steps), temperature, timestep, spacestep, int(restart), 0, None)
        keep = None
        if noise is not None:
            keep = _d(noise)
            assert keep.size == self.n_rep * steps * self.n * 3
            cfg.noise = _ptr(keep, C.c_double)
        self._ck(self.lib.gc_run(self._h, C.byref(cfg)))

    def debug_noise(self, step, replica=0):
        out = np.empty((self.n, 3), dtype=np.float32)
        self._ck(self.lib.gc_debug_noise(self._h, replica, step, _ptr(out, C.c_float)))
        return out

    def compute_force(self, mask=GC_MASK_ALL, replica=0):
        f = np.empty((self.n, 3), dtype=np.float64)
        self._ck(self.lib.gc_compute_force(self._h, replica, mask, _ptr(f, C.c_double)))
        return f

    def compute_energy(self, mask=GC_MASK_ALL, replica=0):
        e = C.c_double()
        self._ck(self.lib.gc_compute_energy(self._h, replica, mask, C.byref(e)))
        return e.value

    def neighbor_pairs(self, radius, replica=0, capacity=None):
        cap = capacity or max(64 * self.n, 1024)
        while True:
            out = np.empty((cap, 2), dtype=np.uint32)
            cnt = C.c_size_t()
            rc = self.lib.gc_neighbor_pairs(self._h, replica, radius, _ptr(out, C.c_uint32), cap,
                                            C.byref(cnt))
            if rc == -5 and capacity is None:
                cap = cnt.value
                continue
            self._ck(rc)
            return out[:cnt.value].copy()

    # ---- contacts ----------------------------------------------------------
    def set_contact_map(self, distance, interval=0, scale_by_core=False, table_capacity=0):
        self._ck(self.lib.gc_set_contact_map(self._h, distance, interval, int(scale_by_core),
                                             table_capacity))

    def contacts_update(self):
        self._ck(self.lib.gc_contacts_update(self._h))

    def contacts_fetch(self, replica=0):
        cnt = C.c_size_t()
        self._ck(self.lib.gc_contacts_fetch(self._h, replica, None, 0, C.byref(cnt)))
        out = np.empty((cnt.value, 3), dtype=np.uint32)
        if cnt.value:
            self._ck(self.lib.gc_contacts_fetch(self._h, replica, _ptr(out, C.c_uint32), cnt.value,
                                                C.byref(cnt)))
        return out

    def contacts_clear(self):
        self._ck(self.lib.gc_contacts_clear(self._h))

    def _device_pairs(self, k, v, n, device):
        """(keys int64[n], counts int32[n]) torch tensors COPIED from context-owned buffers."""
        import torch
        dev = torch.device("cuda", device)
        if n.value == 0:
            return (torch.empty(0, dtype=torch.int64, device=dev),
                    torch.empty(0, dtype=torch.int32, device=dev))
        keys = torch.as_tensor(_DeviceArray(k.value, n.value, "<i8"), device=dev).clone()
        cnts = torch.as_tensor(_DeviceArray(v.value, n.value, "<i4"), device=dev).clone()
        return keys, cnts

    def contacts_reduce_device(self, replica=-1, device=0):
        """Sorted (key = i << 32 | j, count) of one replica, or summed over this
        context's replicas, as CUDA tensors (for an NCCL gather across ranks)."""
        k, v, n = C.c_void_p(), C.c_void_p(), C.c_size_t()
        self._ck(self.lib.gc_contacts_reduce_device(self._h, replica, C.byref(k), C.byref(v),
                                                    C.byref(n)))
        return self._device_pairs(k, v, n, device)

    def contacts_merge_device(self, keys, counts, device=0):
        """Sort + sum (key, count) CUDA tensors gathered from several ranks."""
        keys, counts = keys.contiguous(), counts.contiguous()
        assert keys.dtype.itemsize == 8 and counts.dtype.itemsize == 4 and keys.is_cuda
        k, v, n = C.c_void_p(), C.c_void_p(), C.c_size_t()
        import torch
        torch.cuda.current_stream(keys.device).synchronize()
        self._ck(self.lib.gc_contacts_merge_device(self._h, keys.data_ptr(), counts.data_ptr(),
                                                   keys.numel(), C.byref(k), C.byref(v), C.byref(n)))
        return self._device_pairs(k, v, n, device)

    # ---- spatial decomposition of one large replica (decomp.py drives it) -----
    def decomp_configure(self, rank, world):
        self._decomp_world = int(world)
        self._ck(self.lib.gc_decomp_configure(self._h, int(rank), int(world)))

    def decomp_set_halo(self, layers, capacity=0):
        self._ck(self.lib.gc_decomp_set_halo(self._h, int(layers), int(capacity)))

    def decomp_buffers(self, device=0):
        """dict of flat CUDA tensors viewing the context's exchange buffers: send_lo, send_hi,
        recv_lo, recv_hi (float32 [halo_elems * 4]) and header [16], headers_all [world * 16]
        (float64)."""
        import torch
        p = [C.c_void_p() for _ in range(6)]
        n = C.c_size_t()
        self._ck(self.lib.gc_decomp_buffers(self._h, *[C.byref(x) for x in p], C.byref(n)))
        dev = torch.device("cuda", device)
        out = {}
        for name, x in zip(("send_lo", "send_hi", "recv_lo", "recv_hi"), p[:4]):
            out[name] = torch.as_tensor(_DeviceArray(x.value, n.value * 4, "<f4"), device=dev)
        out["header"] = torch.as_tensor(_DeviceArray(p[4].value, 16, "<f8"), device=dev)
        out["headers_all"] = torch.as_tensor(_DeviceArray(p[5].value, 16 * self._decomp_world, "<f8"),
                                             device=dev)
        return out

    def decomp_step_pack(self):
        self._ck(self.lib.gc_decomp_step_pack(self._h))

    def decomp_set_owned_bound(self, beads):
        self._ck(self.lib.gc_decomp_set_owned_bound(self._h, int(beads)))

    def decomp_recv_buffers(self):
        """Device addresses of recv_lo[0], recv_lo[1], recv_hi[0], recv_hi[1]."""
        out = (C.c_void_p * 4)()
        self._ck(self.lib.gc_decomp_recv_buffers(self._h, out))
        return [out[k] for k in range(4)]

    def decomp_peer_export(self):
        """4 CUDA IPC handles (256 bytes) of this rank's receive buffers."""
        blob = (C.c_ubyte * 256)()
        self._ck(self.lib.gc_decomp_peer_export(self._h, blob))
        return bytes(blob)

    def decomp_peer_attach_ipc(self, lower_blob, upper_blob):
        lo = C.create_string_buffer(lower_blob, 256) if lower_blob else None
        hi = C.create_string_buffer(upper_blob, 256) if upper_blob else None
        self._ck(self.lib.gc_decomp_peer_attach_ipc(self._h, lo, hi))

    def decomp_peer_attach(self, lower=None, upper=None):
        """Same-process neighbours: `lower` / `upper` = their decomp_recv_buffers()."""
        lo = (lower[2], lower[3]) if lower else (None, None)
        hi = (upper[0], upper[1]) if upper else (None, None)
        self._ck(self.lib.gc_decomp_peer_attach(self._h, lo[0], lo[1], hi[0], hi[1]))

    def decomp_peer_detach(self):
        self._ck(self.lib.gc_decomp_peer_detach(self._h))

    def decomp_host_advance(self, steps):
        self._ck(self.lib.gc_decomp_host_advance(self._h, int(steps)))

    def decomp_owned(self):
        """bool [n]: the beads this rank owns now."""
        m = np.zeros(self.n, dtype=np.uint8)
        self._ck(self.lib.gc_decomp_owned(self._h, m.ctypes.data))
        return m.astype(bool)

    def decomp_step_compute(self):
        self._ck(self.lib.gc_decomp_step_compute(self._h))

    def decomp_step_commit(self):
        self._ck(self.lib.gc_decomp_step_commit(self._h))

    # ---- inter-stage transitions on the device (self = destination context) -------
    def transition_interphase_from(self, src, src_ranges, dst_ranges, copy_pairs=(), bits=16,
                                   src_replica=0, dst_replica=0):
        """telophase -> relaxation: spline resampling of every chain of `src` into self."""
        a, b = _i(src_ranges).reshape(-1, 2), _i(dst_ranges).reshape(-1, 2)
        cp = _i(copy_pairs).reshape(-1, 2)
        self._ck(self.lib.gc_transition_interphase(src._h, src_replica, self._h, dst_replica, a.shape[0],
                                                   _ptr(a, C.c_int32), _ptr(b, C.c_int32), cp.shape[0],
                                                   _ptr(cp, C.c_int32) if cp.size else None, int(bits)))

    def transition_prometaphase_from(self, src, src_ranges, target_ranges, sister_ranges, coarse_graining,
                                     displacement, bits=16, src_replica=0, dst_replica=0):
        """interphase -> prometaphase: centroid coarse-graining + displaced sister copy."""
        a, t = _i(src_ranges).reshape(-1, 2), _i(target_ranges).reshape(-1, 2)
        s = _i(sister_ranges).reshape(-1, 2) if sister_ranges is not None else None
        d = _d(displacement)
        self._ck(self.lib.gc_transition_prometaphase(src._h, src_replica, self._h, dst_replica, a.shape[0],
                                                     _ptr(a, C.c_int32), _ptr(t, C.c_int32),
                                                     _ptr(s, C.c_int32) if s is not None else None,
                                                     int(coarse_graining), _ptr(d, C.c_double), int(bits)))

    def transition_cycle_from(self, src, src_ranges, dst_ranges, displacement, bits=16, src_replica=0,
                              dst_replica=0):
        """prometaphase -> next anaphase: target chromatids displaced by -spindle_axis."""
        a, b, d = _i(src_ranges).reshape(-1, 2), _i(dst_ranges).reshape(-1, 2), _d(displacement)
        self._ck(self.lib.gc_transition_cycle(src._h, src_replica, self._h, dst_replica, a.shape[0],
                                              _ptr(a, C.c_int32), _ptr(b, C.c_int32), _ptr(d, C.c_double),
                                              int(bits)))

    # ---- observables / misc -----------------------------------------------
    def radius_of_gyration2(self, ranges, replica=0):
        r = _i(ranges).reshape(-1, 2)
        out = np.empty(r.shape[0], dtype=np.float64)
        self._ck(self.lib.gc_radius_of_gyration(self._h, replica, r.shape[0], _ptr(r, C.c_int32),
                                                _ptr(out, C.c_double)))
        return out

    def radius_of_gyration2_noalloc(self, ranges, replica=0):
        """gc_radius_of_gyration_noalloc: static device scratch, at most 4096 ranges."""
        r = _i(ranges).reshape(-1, 2)
        out = np.empty(r.shape[0], dtype=np.float64)
        self._ck(self.lib.gc_radius_of_gyration_noalloc(self._h, replica, r.shape[0], _ptr(r, C.c_int32),
                                                        _ptr(out, C.c_double)))
        return out

    def debug_grid(self, replica=0):
        out = np.zeros(16)
        self._ck(self.lib.gc_debug_grid(self._h, replica, _ptr(out, C.c_double)))
        return dict(dims=out[:3].astype(int).tolist(), ncell=int(out[3]), origin=out[4:7].tolist(),
                    cell_size=out[7], rc_valid=out[8], max_disp=out[9], need_rebuild=int(out[10]),
                    parity=int(out[11]), dt_cur=out[12])

    def profile_enable(self, enabled=True):
        self._ck(self.lib.gc_profile_enable(self._h, int(enabled)))

    def profile_read(self):
        """{kernel name: (launches, total_ms)} recorded since profile_enable(True)."""
        out = {}
        k = 0
        while True:
            name = C.create_string_buffer(128)
            n, ms = C.c_int64(), C.c_double()
            if self.lib.gc_profile_read(self._h, k, name, 128, C.byref(n), C.byref(ms)) != 0:
                break
            out[name.value.decode()] = (n.value, ms.value)
            k += 1
        return out

    def launch_count(self):
        v = C.c_int64()
        self._ck(self.lib.gc_launch_count(self._h, C.byref(v)))
        return v.value

    def synchronize(self):
        self._ck(self.lib.gc_synchronize(self._h))

    def stream(self):
        v = C.c_void_p()
        self._ck(self.lib.gc_stream(self._h, C.byref(v)))
        return v.value


class Observables:
    """Binned contact matrix + derived observables on one GPU (thin wrapper over gc_obs)."""

    def __init__(self, n_bins, chrom_ranges, device=0):
        self.lib = load_library()
        self.n_bins = int(n_bins)
        r = _i(chrom_ranges).reshape(-1, 2)
        self._h = _P()
        rc = self.lib.gc_obs_create(int(device), self.n_bins, r.shape[0], _ptr(r, C.c_int32), C.byref(self._h))
        if rc != 0:
            msg = self.lib.gc_obs_last_error(self._h).decode() if self._h else "no CUDA device"
            raise GcError(rc, msg)

    def close(self):
        if getattr(self, "_h", None):
            self.lib.gc_obs_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc < 0:
            raise GcError(rc, self.lib.gc_obs_last_error(self._h).decode())

    def set_bead_map(self, bead_bin, bead_chain=None):
        b = _i(bead_bin)
        c = _i(bead_chain) if bead_chain is not None else None
        self._ck(self.lib.gc_obs_set_bead_map(self._h, b.size, _ptr(b, C.c_int32),
                                              _ptr(c, C.c_int32) if c is not None else None))

    def clear(self):
        self._ck(self.lib.gc_obs_clear(self._h))

    def add_runs(self, keys, counts):
        """(keys, counts): CUDA tensors as System.contacts_reduce_device returns them."""
        keys, counts = keys.contiguous(), counts.contiguous()
        self._ck(self.lib.gc_obs_add_runs_device(self._h, keys.data_ptr(), counts.data_ptr(), keys.numel()))

    def add_triples(self, ijv):
        t = np.ascontiguousarray(ijv, dtype=np.uint32).reshape(-1, 3)
        self._ck(self.lib.gc_obs_add_triples(self._h, _ptr(t, C.c_uint32), t.shape[0]))

    def matrix(self):
        out = np.empty((self.n_bins, self.n_bins), dtype=np.float32)
        self._ck(self.lib.gc_obs_get_matrix(self._h, _ptr(out, C.c_float)))
        return out

    def set_matrix(self, m):
        x = np.ascontiguousarray(m, dtype=np.float32).reshape(self.n_bins, self.n_bins)
        self._ck(self.lib.gc_obs_set_matrix(self._h, _ptr(x, C.c_float)))

    def ps_runs(self, keys, counts, max_s):
        keys, counts = keys.contiguous(), counts.contiguous()
        hist = np.zeros(int(max_s) + 1, dtype=np.uint64)
        self._ck(self.lib.gc_obs_ps_runs_device(self._h, keys.data_ptr(), counts.data_ptr(), keys.numel(),
                                                int(max_s), _ptr(hist, C.c_uint64)))
        return hist

    def expected(self):
        n = C.c_int32()
        self._ck(self.lib.gc_obs_max_separation(self._h, C.byref(n)))
        intra, inter = np.empty(n.value), C.c_double()
        self._ck(self.lib.gc_obs_expected(self._h, _ptr(intra, C.c_double), n.value, C.byref(inter)))
        return intra, inter.value

    def pc1(self, mask_intra=False, use_covariance=False, tolerance=1e-4, max_steps=500):
        ev1, pc1 = np.empty(self.n_bins, dtype=np.float32), np.empty(self.n_bins, dtype=np.float32)
        evr, steps = C.c_double(), C.c_int32()
        self._ck(self.lib.gc_obs_pc1(self._h, int(mask_intra), int(use_covariance), float(tolerance),
                                     int(max_steps), _ptr(ev1, C.c_float), _ptr(pc1, C.c_float),
                                     C.byref(evr), C.byref(steps)))
        return dict(ev1=ev1, pc1=pc1, evr=evr.value, steps=steps.value)

    def nci(self):
        out = np.empty(self.n_bins, dtype=np.float32)
        self._ck(self.lib.gc_obs_nci(self._h, _ptr(out, C.c_float)))
        return out

    def type_enrichment(self, bin_type, n_types, min_separation=0):
        t = _i(bin_type)
        out = np.empty((int(n_types), int(n_types)))
        self._ck(self.lib.gc_obs_type_enrichment(self._h, _ptr(t, C.c_int32), int(n_types), int(min_separation),
                                                 _ptr(out, C.c_double)))
        return out
