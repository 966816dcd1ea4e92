"""ctypes loader of the CPU oracle (oracle/maniac_oracle.c).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs,
never by the product package."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from maniac_mc_b200.system import Counters, MmcDesc, MMC_NE, RECORD_DTYPE, System

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libmaniac_oracle.so")


def build(force=False):
    src = os.path.join(_HERE, "maniac_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)
        L.orc_create.restype = C.c_void_p; L.orc_create.argtypes = [C.POINTER(MmcDesc)]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_system_energy.argtypes = [C.c_void_p, dp]
        L.orc_prepare_fourier.argtypes = [C.c_void_p]; L.orc_prepare_fourier.restype = None
        L.orc_num_kvectors.argtypes = [C.c_void_p]
        L.orc_get_kvectors.argtypes = [C.c_void_p, ip, ip, ip, dp, dp]
        L.orc_set_params.argtypes = [C.c_void_p, dp, C.c_double]
        L.orc_set_recalibrate.argtypes = [C.c_void_p, C.c_int]
        L.orc_get_step_sizes.argtypes = [C.c_void_p, dp, dp]
        L.orc_get_state.argtypes = [C.c_void_p, ip, dp, dp]
        L.orc_get_sk.argtypes = [C.c_void_p, dp]
        L.orc_get_reservoir.argtypes = [C.c_void_p, ip, dp]
        L.orc_get_energy.argtypes = [C.c_void_p, dp]
        L.orc_get_counters.argtypes = [C.c_void_p, C.POINTER(Counters)]
        L.orc_pair_singlemol.argtypes = [C.c_void_p, C.c_int, C.c_int, dp, dp]
        L.orc_self_singlemol.restype = C.c_double; L.orc_self_singlemol.argtypes = [C.c_void_p, C.c_int]
        L.orc_intra_singlemol.restype = C.c_double; L.orc_intra_singlemol.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_table_lookup.restype = C.c_double; L.orc_table_lookup.argtypes = [C.c_void_p, C.c_int, C.c_double]
        L.orc_run.argtypes = [C.c_void_p, C.c_int64, dp, C.c_int64, C.c_uint64, C.c_uint32, C.c_uint64, C.c_void_p,
                              C.POINTER(C.c_int64)]
        L.orc_single_move.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, dp, C.c_int64, C.c_void_p]
        L.orc_adjust_step_sizes.argtypes = [C.c_void_p]
        L.orc_ewald_setup.argtypes = [C.c_double, C.c_double, dp, dp, ip, ip]
        L.orc_philox_uniform.restype = C.c_double
        L.orc_philox_uniform.argtypes = [C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint32]
        L.orc_philox.argtypes = [C.POINTER(C.c_uint32), C.c_uint32, C.c_uint32]
        _lib = L
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


class Oracle:
    """One sequential chain, the way the reference runs it."""

    def __init__(self, system: System):
        self.sys = system
        self._desc = system.to_desc()
        self.h = lib().orc_create(C.byref(self._desc))
        self.nres = system.nres

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_destroy(self.h); self.h = None

    def system_energy(self):
        out = np.zeros(MMC_NE); lib().orc_system_energy(self.h, _dp(out)); return out

    def prepare_fourier(self):
        """W(k) + every molecule's phase tables, tracked A(k) and energies zero: the start-up without its O(N K) and O(N^2) sums"""
        lib().orc_prepare_fourier(self.h)

    def energy(self):
        out = np.zeros(MMC_NE); lib().orc_get_energy(self.h, _dp(out)); return out

    def kvectors(self):
        n = lib().orc_num_kvectors(self.h)
        kx, ky, kz = (np.zeros(n, dtype=np.int32) for _ in range(3)); W = np.zeros(n); f = np.zeros(n)
        ip = C.POINTER(C.c_int)
        lib().orc_get_kvectors(self.h, kx.ctypes.data_as(ip), ky.ctypes.data_as(ip), kz.ctypes.data_as(ip), _dp(W), _dp(f))
        return kx, ky, kz, W, f

    def set_params(self, fugacity, temperature):
        f = np.ascontiguousarray(fugacity, dtype=np.float64); lib().orc_set_params(self.h, _dp(f), float(temperature))

    def set_recalibrate(self, flag):
        lib().orc_set_recalibrate(self.h, int(flag))

    def adjust_step_sizes(self):
        lib().orc_adjust_step_sizes(self.h)

    def step_sizes(self):
        a, b = C.c_double(), C.c_double(); lib().orc_get_step_sizes(self.h, C.byref(a), C.byref(b)); return a.value, b.value

    def counts(self):
        c = np.zeros(self.nres, dtype=np.int32)
        lib().orc_get_state(self.h, c.ctypes.data_as(C.POINTER(C.c_int)), None, None); return c

    def state(self):
        c = self.counts()
        nm = int(c.sum()); na = int((c * self.sys.natoms).sum())
        com = np.zeros((max(nm, 1), 3)); off = np.zeros((max(na, 1), 3))
        lib().orc_get_state(self.h, c.ctypes.data_as(C.POINTER(C.c_int)), _dp(com), _dp(off))
        return c, com[:nm], off[:na]

    def reservoir(self):
        """(count[nres], offsets [n, 3]) of the reservoir pool"""
        s = self.sys
        cnt = np.zeros(s.nres, dtype=np.int32)
        lib().orc_get_reservoir(self.h, cnt.ctypes.data_as(C.POINTER(C.c_int)), None)
        off = np.zeros((max(int((cnt * s.natoms).sum()), 1), 3))
        lib().orc_get_reservoir(self.h, cnt.ctypes.data_as(C.POINTER(C.c_int)), _dp(off))
        return cnt, off[:int((cnt * s.natoms).sum())]

    def sk(self):
        n = lib().orc_num_kvectors(self.h); a = np.zeros(2 * n); lib().orc_get_sk(self.h, _dp(a))
        return a[0::2] + 1j * a[1::2]

    def counters(self):
        c = Counters(); lib().orc_get_counters(self.h, C.byref(c)); return c.as_dict()

    def pair_singlemol(self, r, m):
        a, b = C.c_double(), C.c_double(); lib().orc_pair_singlemol(self.h, r, m, C.byref(a), C.byref(b)); return a.value, b.value

    def self_singlemol(self, r):
        return lib().orc_self_singlemol(self.h, r)

    def intra_singlemol(self, r, m):
        return lib().orc_intra_singlemol(self.h, r, m)

    def table_lookup(self, which, r):
        return np.array([lib().orc_table_lookup(self.h, which, float(x)) for x in np.atleast_1d(r)])

    def run(self, n_steps, tape=None, seed=0, chain=0, step0=0, records=True):
        recs = np.zeros(n_steps, dtype=RECORD_DTYPE) if records else None
        used = C.c_int64(0)
        if tape is not None:
            tape = np.ascontiguousarray(tape, dtype=np.float64)
            rc = lib().orc_run(self.h, n_steps, _dp(tape), len(tape), 0, 0, 0, recs.ctypes.data if records else None, C.byref(used))
        else:
            rc = lib().orc_run(self.h, n_steps, None, 0, seed, chain, step0, recs.ctypes.data if records else None, C.byref(used))
        if rc:
            raise RuntimeError(f"oracle run stopped with code {rc} (capacity exceeded?)")
        self.tape_used = used.value
        return recs

    def single_move(self, kind, r, m, tape):
        rec = np.zeros(1, dtype=RECORD_DTYPE)
        tape = np.ascontiguousarray(tape, dtype=np.float64)
        rc = lib().orc_single_move(self.h, kind, r, m, _dp(tape), len(tape), rec.ctypes.data)
        if rc:
            raise RuntimeError(f"oracle single_move failed {rc}")
        return rec[0]


def philox_uniform(seed, chain, step, draw):
    return lib().orc_philox_uniform(seed, chain, step, draw)


def ewald_setup(cutoff, tol, box_len):
    L = np.ascontiguousarray(box_len, dtype=np.float64); a = C.c_double(); k = (C.c_int * 3)(); n = C.c_int()
    lib().orc_ewald_setup(cutoff, tol, _dp(L), C.byref(a), k, C.byref(n))
    return a.value, list(k), n.value
