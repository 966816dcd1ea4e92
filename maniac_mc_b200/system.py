"""Packed system description (the contents of ``mmc_desc``, include/maniac_gpu.h) and builders.

``System`` is the host-side mirror of the parts of ``simulation_state`` (simulation_state.f90)
that the per-move energy path reads: residue table, charges/types, LJ table compacted by atom
type, packed COM/offset coordinates, orthorhombic box, Ewald parameters, move parameters.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, field

import numpy as np

from . import readers

MMC_NE = 6
MMC_CREATION, MMC_DELETION, MMC_TRANSLATION, MMC_ROTATION, MMC_SWAP = 1, 2, 3, 4, 5
E_NONCOUL, E_COUL, E_RECIP, E_SELF, E_INTRA, E_TOTAL = range(6)


class MmcDesc(C.Structure):
    _fields_ = [
        ("nres", C.c_int), ("max_atoms", C.c_int), ("ntypes", C.c_int), ("n_chains", C.c_int),
        ("natoms", C.POINTER(C.c_int)), ("is_active", C.POINTER(C.c_int)),
        ("count", C.POINTER(C.c_int)), ("capacity", C.POINTER(C.c_int)),
        ("charges", C.POINTER(C.c_double)), ("types", C.POINTER(C.c_int)),
        ("sigma", C.POINTER(C.c_double)), ("epsilon", C.POINTER(C.c_double)),
        ("com", C.POINTER(C.c_double)), ("offsets", C.POINTER(C.c_double)),
        ("box_lo", C.c_double * 3), ("box_len", C.c_double * 3),
        ("alpha", C.c_double), ("cutoff", C.c_double), ("temperature", C.c_double),
        ("fugacity", C.POINTER(C.c_double)),
        ("kmax", C.c_int * 3),
        ("translation_step", C.c_double), ("rotation_step", C.c_double),
        ("p_translation", C.c_double), ("p_rotation", C.c_double),
        ("init_sk", C.c_int), ("deletion_mode", C.c_int), ("use_table", C.c_int), ("device", C.c_int),
        ("coulomb_xcut", C.c_double),
        ("shard_rank", C.c_int),
        ("shard_count", C.c_int),
        ("p_swap", C.c_double),
        ("reservoir_count", C.POINTER(C.c_int)), ("reservoir_offsets", C.POINTER(C.c_double)),
        ("tilt", C.c_double * 3),
    ]


class MoveRecord(C.Structure):
    _fields_ = [("kind", C.c_int32), ("res", C.c_int32), ("mol", C.c_int32), ("accepted", C.c_int32),
                ("old_e", C.c_double * MMC_NE), ("new_e", C.c_double * MMC_NE), ("prob", C.c_double)]


class Counters(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("translations", "rotations", "creations", "deletions",
                                         "trial_translations", "trial_rotations", "trial_creations", "trial_deletions",
                                         "swaps", "trial_swaps")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


N_COUNTERS = len(Counters._fields_)   # int64 words of one mmc_counters

RECORD_DTYPE = np.dtype([("kind", "<i4"), ("res", "<i4"), ("mol", "<i4"), ("accepted", "<i4"),
                         ("old_e", "<f8", (MMC_NE,)), ("new_e", "<f8", (MMC_NE,)), ("prob", "<f8")])
assert RECORD_DTYPE.itemsize == C.sizeof(MoveRecord)


def ewald_setup(cutoff, tol, box_len):
    """ComputeEwaldParameters / ComputeFourierIndices, prepare_utils.f90:169-214 (host set-up)."""
    tol = min(abs(tol), 0.5)
    s = math.sqrt(abs(math.log(tol * cutoff)))
    alpha = math.sqrt(abs(math.log(tol * cutoff * s))) / cutoff
    p = math.sqrt(-math.log(tol * cutoff * (2.0 * s * alpha) ** 2))
    kmax = [int(math.floor(0.25 + L * alpha * p / math.pi + 0.5)) for L in box_len]
    return alpha, kmax


def count_kvectors(kmax):
    kx = np.arange(0, kmax[0] + 1)[:, None, None] / kmax[0]
    ky = np.arange(-kmax[1], kmax[1] + 1)[None, :, None] / kmax[1]
    kz = np.arange(-kmax[2], kmax[2] + 1)[None, None, :] / kmax[2]
    k2 = kx ** 2 + ky ** 2 + kz ** 2
    return int(((np.abs(k2) >= 1e-10) & (k2 <= 1.0)).sum())


@dataclass
class System:
    natoms: np.ndarray        # [nres] int32
    is_active: np.ndarray     # [nres] int32
    count: np.ndarray         # [nres] int32
    capacity: np.ndarray      # [nres] int32
    charges: np.ndarray       # [nres, max_atoms] f8
    types: np.ndarray         # [nres, max_atoms] int32 (1-based)
    sigma: np.ndarray         # [ntypes, ntypes]
    epsilon: np.ndarray       # [ntypes, ntypes] Kelvin
    com: np.ndarray           # [sum(count), 3]
    offsets: np.ndarray       # [sum(count*natoms), 3]
    box_lo: np.ndarray
    box_len: np.ndarray
    alpha: float
    cutoff: float
    temperature: float
    fugacity: np.ndarray      # [nres] 1/A^3
    kmax: list
    translation_step: float = 1.0
    rotation_step: float = 0.685
    p_translation: float = 0.25
    p_rotation: float = 0.25
    p_swap: float = 0.0         # proba%swap; 0 = the reference's de-facto behaviour (the move is never called there)
    init_sk: int = 0
    deletion_mode: int = 0
    use_table: int = 0
    coulomb_xcut: float = 0.0   # 0 = library default (6.0), negative = evaluate every pair
    names: list = field(default_factory=list)
    reservoir_count: np.ndarray | None = None     # [nres] molecules in the reservoir box per type (None = no reservoir)
    reservoir_offsets: np.ndarray | None = None   # [sum(count*natoms), 3] their site offsets, packed like `offsets`
    ewald_tolerance: float = 1e-5
    tilt: np.ndarray = field(default_factory=lambda: np.zeros(3))   # xy, xz, yz (box%tilt); zeros = orthorhombic

    # ---- helpers ----
    @property
    def nres(self):
        return len(self.natoms)

    @property
    def volume(self):
        return float(np.prod(self.box_len))

    @property
    def num_kvectors(self):
        return count_kvectors(self.kmax)

    def total_atoms(self):
        return int((self.count * self.natoms).sum())

    def mol_offsets(self):
        """start index of each residue in the packed com / offsets arrays"""
        mo = np.concatenate([[0], np.cumsum(self.count)])[:-1]
        ao = np.concatenate([[0], np.cumsum(self.count * self.natoms)])[:-1]
        return mo.astype(int), ao.astype(int)

    def positions(self):
        """absolute atom positions com + offset in packed order"""
        mo, ao = self.mol_offsets()
        out = np.empty_like(self.offsets)
        for r in range(self.nres):
            n, na = int(self.count[r]), int(self.natoms[r])
            c = np.repeat(self.com[mo[r]:mo[r] + n], na, axis=0)
            out[ao[r]:ao[r] + n * na] = c + self.offsets[ao[r]:ao[r] + n * na]
        return out

    def to_desc(self, n_chains=1, device=0, shard_rank=0, shard_count=1):
        """ctypes mmc_desc; keeps the backing arrays alive on the returned object."""
        d = MmcDesc()
        keep = {}

        def arr(name, a, dt):
            a = np.ascontiguousarray(a, dtype=dt)
            keep[name] = a
            ct = C.c_int if dt == np.int32 else C.c_double
            return a.ctypes.data_as(C.POINTER(ct))

        d.nres, d.max_atoms, d.ntypes, d.n_chains = self.nres, self.charges.shape[1], self.sigma.shape[0], n_chains
        d.natoms = arr("natoms", self.natoms, np.int32)
        d.is_active = arr("is_active", self.is_active, np.int32)
        d.count = arr("count", self.count, np.int32)
        d.capacity = arr("capacity", self.capacity, np.int32)
        d.charges = arr("charges", self.charges, np.float64)
        d.types = arr("types", self.types, np.int32)
        d.sigma = arr("sigma", self.sigma, np.float64)
        d.epsilon = arr("epsilon", self.epsilon, np.float64)
        d.com = arr("com", self.com.reshape(-1) if self.com.size else np.zeros(3), np.float64)
        d.offsets = arr("offsets", self.offsets.reshape(-1) if self.offsets.size else np.zeros(3), np.float64)
        d.fugacity = arr("fugacity", self.fugacity, np.float64)
        for k in range(3):
            d.box_lo[k], d.box_len[k], d.kmax[k] = float(self.box_lo[k]), float(self.box_len[k]), int(self.kmax[k])
        d.alpha, d.cutoff, d.temperature = self.alpha, self.cutoff, self.temperature
        d.translation_step, d.rotation_step = self.translation_step, self.rotation_step
        d.p_translation, d.p_rotation = self.p_translation, self.p_rotation
        d.p_swap = float(self.p_swap)
        d.init_sk, d.deletion_mode, d.use_table, d.device = self.init_sk, self.deletion_mode, self.use_table, device
        d.coulomb_xcut = float(self.coulomb_xcut)
        d.shard_rank, d.shard_count = int(shard_rank), int(shard_count)
        for k in range(3):
            d.tilt[k] = float(self.tilt[k])
        if self.reservoir_count is not None:
            d.reservoir_count = arr("reservoir_count", self.reservoir_count, np.int32)
            ro = np.asarray(self.reservoir_offsets, dtype=np.float64)
            d.reservoir_offsets = arr("reservoir_offsets", ro.reshape(-1) if ro.size else np.zeros(3), np.float64)
        d._keep = keep
        return d

    # ---- persistence (tests/golden fixtures) ----
    _ARRS = ("natoms", "is_active", "count", "capacity", "charges", "types", "sigma", "epsilon", "com",
             "offsets", "box_lo", "box_len", "fugacity")
    _SCAL = ("alpha", "cutoff", "temperature", "translation_step", "rotation_step", "p_translation",
             "p_rotation", "p_swap", "init_sk", "deletion_mode", "use_table", "ewald_tolerance")

    def save(self, path):
        d = {k: getattr(self, k) for k in self._ARRS}
        d.update({k: np.array(getattr(self, k)) for k in self._SCAL})
        d["kmax"] = np.array(self.kmax, dtype=np.int32)
        if np.any(np.asarray(self.tilt) != 0.0):
            d["tilt"] = np.asarray(self.tilt, dtype=np.float64)
        np.savez_compressed(path, **d)

    @classmethod
    def load(cls, path):
        z = np.load(path)
        kw = {k: z[k] for k in cls._ARRS}
        for k in cls._SCAL:
            if k not in z.files:
                continue
            v = z[k].item()
            kw[k] = int(v) if k in ("init_sk", "deletion_mode", "use_table") else float(v)
        kw["kmax"] = [int(v) for v in z["kmax"]]
        if "tilt" in z.files:
            kw["tilt"] = z["tilt"]
        return cls(**kw)

    def copy(self):
        import copy
        return copy.deepcopy(self)
