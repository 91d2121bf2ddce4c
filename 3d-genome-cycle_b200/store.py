"""Trajectory / contact output path with the reference's on-disk schema.

The reference keeps everything in one HDF5 file written by `simulation_store`
(SIM/common/simulation_store.{hpp,cpp}) and `prepare init` (SIM/stage_prepare/prepare.cpp:374-562):

  /metadata/{master_seed u32, config str, config_source str, chains_source str}
  /stages/<stage>/metadata/{seed, particle_types i32 enum, ab_factors f32 (N,2), chain_names,
                            chain_ranges i32 (C,2), nucleolar_bonds i32 (B,2), kinetochore_beads,
                            sister_chromatids i32 (C,2), pole_positions f32 (2,3)}
  /stages/<stage>/.steps                      decimal STRINGS, one per saved frame
  /stages/<stage>/<step>/positions            f32 (N,3), 16 significant bits, shuffle + deflate 6
  /stages/<stage>/<step>/context              str (JSON: time, wall_semiaxes, core_scale,
                                              bond_scale, mean_energy, wall_energy)
  /stages/<stage>/<step>/contacts             i32 (M,3) sorted (i, j, count), deflate 4

libhdf5 / h5py are not installed in this image, so the same hierarchy is written as a
directory tree: every dataset path becomes `<root>/<path>.npy` (arrays) or `<root>/<path>.json`
(strings and string lists) -- `FlatStore`.  When h5py is importable, `H5Store` writes the real
thing with the reference's dataset names, dtypes and filters.  Both expose the methods of the
reference class with the same names and argument meaning (simulation_store.hpp:65-111), so a
re-hosted driver reads like the reference's.

Frames arrive ALREADY quantised from the device (gc_get_frame / gc_frame_begin + gc_frame_end do
the 16-bit rounding of simulation_store.cpp:22-32 in a kernel and the copy overlaps the next
steps); save_positions therefore stores f32 as is unless told to quantise host data.
"""
from __future__ import annotations

import json
import shutil
from pathlib import Path

import numpy as np

from .synth import quantize_f32

INTERPHASE_TYPES = {"unknown": 0, "a": 1, "b": 2, "u": 3, "centromere": 4, "active_nor": 5,
                    "silent_nor": 6, "nucleolus": 7}
_TAG_TO_TYPE = {"A": "a", "B": "b", "u": "u", "cen": "centromere", "anor": "active_nor",
                "bnor": "silent_nor", "nucleolus": "nucleolus", "unknown": "unknown"}
MITOTIC_TYPES = {"unknown": 0, "arm": 1, "kinetochore": 2}


class StoreError(RuntimeError):
    """simulation_store_error (simulation_store.hpp:114-118)."""


def interphase_context_json(ctx, mean_energy=0.0, wall_energy=0.0) -> str:
    """jsoncons::encode_json(interphase_context) (simulation_store.cpp:36-46): member order
    time, wall_semiaxes, core_scale, bond_scale, mean_energy, wall_energy."""
    return json.dumps({"time": float(ctx.time), "wall_semiaxes": [float(v) for v in ctx.semiaxes],
                       "core_scale": float(ctx.core_scale), "bond_scale": float(ctx.bond_scale),
                       "mean_energy": float(mean_energy), "wall_energy": float(wall_energy)},
                      separators=(",", ":"))


class _Base:
    def __init__(self):
        self._stage = "unknown"

    # -- hierarchy helpers (simulation_store.cpp: locate_*) ---------------------------------
    def set_stage(self, name: str):
        self._stage = name

    def _data(self, *parts) -> str:
        return "/".join(["/stages", self._stage] + [str(p) for p in parts])

    def _meta(self, stage, key) -> str:
        return f"/stages/{stage}/metadata/{key}"

    # -- snapshots ---------------------------------------------------------------------------
    def clear_frames(self):
        if self._exists(self._data(".steps")):
            self._write_strings(self._data(".steps"), [])

    def append_frame(self, step: int):
        idx = self._read_strings(self._data(".steps")) if self._exists(self._data(".steps")) else []
        idx.append(str(int(step)))
        self._write_strings(self._data(".steps"), idx)

    def load_steps(self):
        if not self._exists(self._data(".steps")):
            return []
        return [int(s) for s in self._read_strings(self._data(".steps"))]

    def save_positions(self, step: int, positions, quantize=False):
        x = np.asarray(positions)
        if x.ndim != 2 or x.shape[1] != 3:
            raise StoreError("positions must be (N, 3)")
        x = quantize_f32(x) if quantize else np.ascontiguousarray(x, dtype=np.float32)
        self._write_array(self._data(int(step), "positions"), x, compression=6, shuffle=True)

    def load_positions(self, step: int):
        path = self._data(int(step), "positions")
        if not self._exists(path):
            raise StoreError(f"no positions at {path}")
        return self._read_array(path).astype(np.float64)

    def check_positions(self, step: int) -> bool:
        return self._exists(self._data(int(step), "positions"))

    def save_interphase_context(self, step: int, ctx, mean_energy=0.0, wall_energy=0.0):
        self._write_string(self._data(int(step), "context"), interphase_context_json(ctx, mean_energy, wall_energy))

    def load_interphase_context(self, step: int) -> dict:
        return json.loads(self._read_string(self._data(int(step), "context")))

    def save_contacts(self, step: int, contacts):
        c = np.asarray(contacts)
        if c.size == 0:
            return                                   # simulation_store.cpp:257-259
        self._write_array(self._data(int(step), "contacts"), c.reshape(-1, 3).astype(np.int32), compression=4)

    def load_contacts(self, step: int):
        return self._read_array(self._data(int(step), "contacts"))

    # -- metadata (prepare.cpp:374-562) -----------------------------------------------------------
    def write_prepared(self, prepared, config_json="{}", chains_source="", config_source="", master_seed=0):
        self._write_array("/metadata/master_seed", np.asarray(master_seed, dtype=np.uint32))
        self._write_string("/metadata/config", config_json)
        self._write_string("/metadata/config_source", config_source)
        self._write_string("/metadata/chains_source", chains_source)
        i, a, p = prepared.interphase, prepared.anatelophase, prepared.prometaphase
        types = np.asarray([INTERPHASE_TYPES[_TAG_TO_TYPE[t]] for t in prepared.particle_types], dtype=np.int32)
        for stage in ("interphase",):
            self._write_array(self._meta(stage, "seed"), np.asarray(i.seed, dtype=np.uint32))
            self._write_array(self._meta(stage, "particle_types"), types)
            self._write_array(self._meta(stage, "ab_factors"), np.stack([i.a, i.b], axis=1).astype(np.float32))
            self._write_strings(self._meta(stage, "chain_names"), list(prepared.chain_names))
            self._write_array(self._meta(stage, "chain_ranges"), np.asarray(i.chains, dtype=np.int32))
            self._write_array(self._meta(stage, "nucleolar_bonds"), np.asarray(i.nucleolar_bonds, dtype=np.int32))
        # relaxation shares the interphase metadata (soft links in the reference, h5_misc.hpp:9-27)
        self._link("/stages/relaxation/metadata", "/stages/interphase/metadata")
        mt = np.asarray([MITOTIC_TYPES[t] for t in prepared.mitotic_types], dtype=np.int32)
        self._write_array(self._meta("anaphase", "seed"), np.asarray(a.seed, dtype=np.uint32))
        self._write_array(self._meta("anaphase", "particle_types"), mt)
        self._write_strings(self._meta("anaphase", "chain_names"), list(prepared.chain_names))
        self._write_array(self._meta("anaphase", "chain_ranges"), np.asarray(a.chains, dtype=np.int32))
        self._write_array(self._meta("anaphase", "kinetochore_beads"), np.asarray(a.kinetochores, dtype=np.int32))
        self._link("/stages/telophase/metadata", "/stages/anaphase/metadata")
        pt = np.concatenate([np.tile(mt[s:e], 2) for s, e in a.chains]) if len(a.chains) else mt
        self._write_array(self._meta("prometaphase", "seed"), np.asarray(p.seed, dtype=np.uint32))
        self._write_array(self._meta("prometaphase", "particle_types"), pt.astype(np.int32))
        self._write_strings(self._meta("prometaphase", "chain_names"), list(prepared.prometaphase_chain_names))
        self._write_array(self._meta("prometaphase", "chain_ranges"), np.asarray(p.chains, dtype=np.int32))
        self._write_array(self._meta("prometaphase", "kinetochore_beads"), np.asarray(p.kinetochores, dtype=np.int32))
        self._write_array(self._meta("prometaphase", "sister_chromatids"), np.asarray(p.sisters, dtype=np.int32))
        self._write_array(self._meta("prometaphase", "pole_positions"), np.asarray(p.poles, dtype=np.float32))

    def load_seed(self, stage) -> int:
        return int(self._read_array(self._meta(stage, "seed")))

    def load_chains(self, stage):
        return self._read_array(self._meta(stage, "chain_ranges"))


class FlatStore(_Base):
    """The schema above as a directory tree (no HDF5 in this image)."""

    def __init__(self, root, create=True):
        super().__init__()
        self.root = Path(root)
        if create:
            self.root.mkdir(parents=True, exist_ok=True)
        elif not self.root.is_dir():
            raise StoreError(f"no store at {self.root}")
        self._links = {"/" + str(f.relative_to(self.root))[:-5]: f.read_text().strip()
                       for f in self.root.rglob("*.link")}

    def _p(self, path: str, ext: str) -> Path:
        for src, target in self._links.items():       # soft links of the metadata groups
            if path == src or path.startswith(src + "/"):
                path = target + path[len(src):]
                break
        p = self.root / path.lstrip("/")
        return p.with_name(p.name + ext)

    def _exists(self, path):
        return self._p(path, ".npy").exists() or self._p(path, ".json").exists()

    def _write_array(self, path, arr, compression=0, shuffle=False):
        f = self._p(path, ".npy")
        f.parent.mkdir(parents=True, exist_ok=True)
        np.save(f, np.asarray(arr))

    def _read_array(self, path):
        f = self._p(path, ".npy")
        if not f.exists():
            raise StoreError(f"no dataset {path}")
        return np.load(f)

    def _write_string(self, path, text):
        f = self._p(path, ".json")
        f.parent.mkdir(parents=True, exist_ok=True)
        f.write_text(json.dumps(text))

    def _read_string(self, path):
        f = self._p(path, ".json")
        if not f.exists():
            raise StoreError(f"no dataset {path}")
        return json.loads(f.read_text())

    _write_strings = _write_string
    _read_strings = _read_string

    def _link(self, path, target):
        f = self.root / (path.lstrip("/") + ".link")
        f.parent.mkdir(parents=True, exist_ok=True)
        f.write_text(target)
        self._links[path] = target

    def remove(self):
        shutil.rmtree(self.root, ignore_errors=True)


class H5Store(_Base):
    """The same through h5py (dataset names, dtypes and filters of the reference)."""

    def __init__(self, filename, mode="a"):
        super().__init__()
        try:
            import h5py
        except ImportError as ex:          # this image: fall back to FlatStore
            raise StoreError("h5py is not installed: use FlatStore") from ex
        self._h5py = h5py
        self.f = h5py.File(filename, mode)

    def _exists(self, path):
        return path in self.f

    def _write_array(self, path, arr, compression=0, shuffle=False):
        if path in self.f:
            del self.f[path]
        arr = np.asarray(arr)
        kw = {}
        if compression and arr.ndim > 0 and arr.size > 0:
            kw = {"compression": "gzip", "compression_opts": int(compression), "shuffle": bool(shuffle)}
        self.f.create_dataset(path, data=arr, **kw)

    def _read_array(self, path):
        if path not in self.f:
            raise StoreError(f"no dataset {path}")
        return self.f[path][()]

    def _write_string(self, path, text):
        if path in self.f:
            del self.f[path]
        self.f.create_dataset(path, data=text, dtype=self._h5py.string_dtype())

    def _read_string(self, path):
        v = self._read_array(path)
        return v.decode() if isinstance(v, bytes) else str(v)

    def _write_strings(self, path, items):
        if path in self.f:
            del self.f[path]
        self.f.create_dataset(path, data=np.asarray(items, dtype=object), dtype=self._h5py.string_dtype(),
                              shape=(len(items),))

    def _read_strings(self, path):
        return [v.decode() if isinstance(v, bytes) else str(v) for v in self._read_array(path)]

    def _link(self, path, target):
        if path in self.f:
            del self.f[path]
        self.f[path] = self._h5py.SoftLink(target)

    def close(self):
        self.f.close()


def open_store(path):
    """H5Store when h5py exists and the path ends in .h5, else FlatStore."""
    if str(path).endswith(".h5"):
        try:
            return H5Store(path)
        except StoreError:
            pass
    return FlatStore(path)


class TrajectoryWriter:
    """The output half of a driver's sampling callback
    (SIM/stage_interphase/simulation_driver_interphase.cpp:29-49): frame -> save_positions,
    save_interphase_context, append_frame; every `output_window`-th frame also the contact
    triples.  Usable as run_cell_cycle(frame_sink=...)."""

    def __init__(self, store, sampling_interval=1000):
        self.store = store
        self.interval = int(sampling_interval)
        self.frame_no = {}

    def begin_stage(self, name):
        self.store.set_stage(name)
        self.store.clear_frames()
        self.frame_no[name] = 0
        self.stage = name

    def __call__(self, system, frame, step=None, contacts=None, mean_energy=0.0):
        k = self.frame_no.get(self.stage, 0) + 1
        self.frame_no[self.stage] = k
        step = int(step) if step is not None else k * self.interval
        self.store.save_positions(step, np.asarray(frame).reshape(-1, 3))
        if self.stage in ("interphase", "relaxation"):
            self.store.save_interphase_context(step, system.get_context(0), mean_energy=mean_energy)
        if contacts is not None:
            self.store.save_contacts(step, contacts)
        self.store.append_frame(step)
