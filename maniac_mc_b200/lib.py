"""ctypes binding of libmaniacgpu.so (include/maniac_gpu.h).

This is harness glue: the product is the shared library and its C ABI.  There is no CPU
fallback anywhere: if the library is missing or no CUDA device is present, calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from .system import Counters, MmcDesc, MMC_NE, MoveRecord, RECORD_DTYPE, System

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
SO_PATH = os.environ.get("MANIAC_GPU_LIB") or os.path.join(_PKG, "libmaniacgpu.so")   # env override: A/B builds side by side
SOURCES = [os.path.join(_PKG, "csrc", f) for f in ("maniac_gpu.cu", "common.cuh", "kernels_move.cuh", "kernels_sweep.cuh", "kernels_chain.cuh",
                                                        "kernels_system.cuh")] + \
          [os.path.join(_ROOT, "include", "maniac_gpu.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


class ManiacGpuError(RuntimeError):
    pass


def build(force=False, verbose=False):
    """nvcc cross-compiles for sm_100a (works without a GPU)."""
    newest = max(os.path.getmtime(p) for p in SOURCES)
    if force or not os.path.exists(SO_PATH) or os.path.getmtime(SO_PATH) < newest:
        nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
        cmd = [nvcc] + NVCC_FLAGS + ["-o", SO_PATH, SOURCES[0]]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return SO_PATH


# every symbol declared in include/maniac_gpu.h
ABI_SYMBOLS = [
    "mmc_ewald_setup", "mmc_init", "mmc_destroy", "mmc_last_error", "mmc_num_kvectors", "mmc_get_kvectors",
    "mmc_set_chain_params", "mmc_set_chain_id", "mmc_adjust_step_sizes", "mmc_get_step_sizes", "mmc_set_chain_state", "mmc_get_chain_state", "mmc_set_step_sizes",
    "mmc_system_energy", "mmc_trial_energy", "mmc_trial_energy_swap", "mmc_trial_commit", "mmc_sync_to_host", "mmc_get_sk", "mmc_get_reservoir",
    "mmc_run_batch", "mmc_run_batch_async", "mmc_synchronize", "mmc_tape_used",
    "mmc_dbg_pair_singlemol", "mmc_dbg_self_singlemol", "mmc_dbg_intra_singlemol", "mmc_dbg_table_lookup",
    "mmc_dbg_erfc", "mmc_stream", "mmc_launch_count", "mmc_measure_fp64_peak",
    "mmc_chain_capacity", "mmc_upload_chains", "mmc_download_chains", "mmc_profile_kernels", "mmc_pair_work",
    "mmc_shard_step_begin", "mmc_shard_step_end", "mmc_shard_trial_begin", "mmc_shard_trial_end", "mmc_shard_range",
    "mmc_snapshot_begin", "mmc_snapshot_wait", "mmc_dbg_set_tracked", "mmc_system_energy_all", "mmc_dbg_phase_cycles", "mmc_dbg_fw_cache", "mmc_trial_step", "mmc_dbg_strict_loop", "mmc_erfc_noalloc", "mmc_chain_launch_config",
]

_lib = None


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ManiacGpuError(f"{SO_PATH} not built; run `python -c 'import __graft_entry__ as g; g.build()'` "
                             "(there is no CPU fallback)")
    L = C.CDLL(SO_PATH)
    dp, ip, vp = C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_void_p
    L.mmc_last_error.restype = C.c_char_p
    L.mmc_ewald_setup.argtypes = [C.c_double, C.c_double, dp, dp, ip, ip]
    L.mmc_init.argtypes = [C.POINTER(vp), C.POINTER(MmcDesc)]
    L.mmc_destroy.argtypes = [vp]; L.mmc_destroy.restype = None
    L.mmc_num_kvectors.argtypes = [vp]
    L.mmc_get_kvectors.argtypes = [vp, ip, ip, ip, dp, dp]
    L.mmc_set_chain_params.argtypes = [vp, C.c_int, dp, C.c_double]
    L.mmc_set_chain_state.argtypes = [vp, C.c_int, ip, dp, dp]
    L.mmc_set_chain_id.argtypes = [vp, C.c_int, C.c_uint32]
    L.mmc_adjust_step_sizes.argtypes = [vp]
    L.mmc_get_step_sizes.argtypes = [vp, C.c_int, dp, dp]
    L.mmc_get_chain_state.argtypes = [vp, C.c_int, ip, dp, dp]
    L.mmc_set_step_sizes.argtypes = [vp, C.c_int, C.c_double, C.c_double]
    L.mmc_system_energy.argtypes = [vp, C.c_int, dp]
    L.mmc_system_energy_all.argtypes = [vp, dp]
    L.mmc_trial_energy.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, dp, dp, dp, dp]
    L.mmc_trial_commit.argtypes = [vp, C.c_int, C.c_int]
    L.mmc_trial_step.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, dp, dp, C.c_double, dp, dp, dp, ip]
    L.mmc_trial_energy_swap.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, dp, dp, dp, dp]
    L.mmc_sync_to_host.argtypes = [vp, C.c_int, ip, dp, dp, dp]
    L.mmc_get_sk.argtypes = [vp, C.c_int, dp]
    L.mmc_get_reservoir.argtypes = [vp, C.c_int, ip, dp]
    L.mmc_run_batch.argtypes = [vp, C.c_int64, C.c_uint64, dp, C.c_int64, vp, vp]
    L.mmc_run_batch_async.argtypes = [vp, C.c_int64, C.c_uint64]
    L.mmc_synchronize.argtypes = [vp]
    L.mmc_tape_used.argtypes = [vp, C.c_int]; L.mmc_tape_used.restype = C.c_int64
    L.mmc_dbg_pair_singlemol.argtypes = [vp, C.c_int, C.c_int, C.c_int, dp, dp]
    L.mmc_dbg_self_singlemol.argtypes = [vp, C.c_int, dp]
    L.mmc_dbg_intra_singlemol.argtypes = [vp, C.c_int, C.c_int, C.c_int, dp]
    L.mmc_dbg_table_lookup.argtypes = [vp, C.c_int, dp, C.c_int, dp]
    L.mmc_dbg_erfc.argtypes = [vp, dp, C.c_int, dp]
    L.mmc_erfc_noalloc.argtypes = [vp, dp, C.c_int, dp]
    L.mmc_dbg_set_tracked.argtypes = [vp, C.c_int, dp, dp]
    L.mmc_dbg_strict_loop.argtypes = [vp, C.c_int, C.c_int, C.c_uint64, C.c_int, dp, C.POINTER(C.c_int64)]
    L.mmc_dbg_phase_cycles.argtypes = [vp, C.POINTER(C.c_int64)]
    L.mmc_dbg_fw_cache.argtypes = [vp, C.c_int, dp]
    L.mmc_stream.argtypes = [vp]; L.mmc_stream.restype = vp
    L.mmc_launch_count.argtypes = [vp]; L.mmc_launch_count.restype = C.c_int64
    L.mmc_chain_launch_config.argtypes = [vp, ip, ip]
    L.mmc_pair_work.argtypes = [vp, C.POINTER(C.c_int64)]
    L.mmc_shard_step_begin.argtypes = [vp, C.c_uint64, vp]
    L.mmc_shard_step_end.argtypes = [vp, vp, vp]
    L.mmc_shard_trial_begin.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, dp, dp, vp]
    L.mmc_shard_trial_end.argtypes = [vp, C.c_int, vp, dp, dp]
    L.mmc_shard_range.argtypes = [vp, ip, ip]
    L.mmc_measure_fp64_peak.argtypes = [C.c_int, dp]
    L.mmc_chain_capacity.argtypes = [vp, ip, ip]
    L.mmc_upload_chains.argtypes = [vp, vp, vp, vp]
    L.mmc_download_chains.argtypes = [vp, vp, vp, vp, vp, vp]
    L.mmc_snapshot_begin.argtypes = [vp, C.c_int]
    L.mmc_snapshot_wait.argtypes = [vp, C.c_int] + [C.POINTER(vp)] * 5
    L.mmc_profile_kernels.argtypes = [vp, C.c_int, C.c_uint64, dp, C.POINTER(C.c_int64)]
    _lib = L
    return L


def _dp(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_int))


def _addr(a):
    if a is None:
        return None
    if isinstance(a, int):
        return a
    return a.ctypes.data


def _ck(rc):
    if rc != 0:
        raise ManiacGpuError(load().mmc_last_error().decode())


def ewald_setup(cutoff, tol, box_len):
    L = np.ascontiguousarray(box_len, dtype=np.float64)
    a, k, n = C.c_double(), (C.c_int * 3)(), C.c_int()
    _ck(load().mmc_ewald_setup(cutoff, tol, _dp(L), C.byref(a), k, C.byref(n)))
    return a.value, list(k), n.value


def measure_fp64_peak(device=0):
    t = C.c_double()
    _ck(load().mmc_measure_fp64_peak(device, C.byref(t)))
    return t.value


class ManiacGPU:
    """Device-resident state of n_chains chains of one system."""

    def __init__(self, system: System, n_chains=1, device=0, shard_rank=0, shard_count=1):
        self.sys = system
        self.n_chains = n_chains
        self._desc = system.to_desc(n_chains=n_chains, device=device, shard_rank=shard_rank, shard_count=shard_count)
        self.h = C.c_void_p()
        _ck(load().mmc_init(C.byref(self.h), C.byref(self._desc)))
        self.nres = system.nres
        self.nk = load().mmc_num_kvectors(self.h)

    def close(self):
        if getattr(self, "h", None):
            load().mmc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- set-up
    def kvectors(self):
        n = self.nk
        kx, ky, kz = (np.zeros(n, dtype=np.int32) for _ in range(3)); W = np.zeros(n); f = np.zeros(n)
        _ck(load().mmc_get_kvectors(self.h, _ip(kx), _ip(ky), _ip(kz), _dp(W), _dp(f)))
        return kx, ky, kz, W, f

    def set_chain_params(self, chain, fugacity, temperature):
        f = np.ascontiguousarray(fugacity, dtype=np.float64)
        _ck(load().mmc_set_chain_params(self.h, chain, _dp(f), float(temperature)))

    def set_chain_id(self, chain, stream_id):
        _ck(load().mmc_set_chain_id(self.h, chain, int(stream_id)))

    def adjust_step_sizes(self):
        """AdjustMoveStepSizes for every chain (once per block)"""
        _ck(load().mmc_adjust_step_sizes(self.h))

    def step_sizes(self, chain=0):
        a, b = C.c_double(), C.c_double()
        _ck(load().mmc_get_step_sizes(self.h, chain, C.byref(a), C.byref(b)))
        return a.value, b.value

    def set_step_sizes(self, chain, tstep, rstep):
        _ck(load().mmc_set_step_sizes(self.h, chain, float(tstep), float(rstep)))

    def set_chain_state(self, chain, count, com, offsets):
        cnt = np.ascontiguousarray(count, dtype=np.int32)
        com = np.ascontiguousarray(com, dtype=np.float64); off = np.ascontiguousarray(offsets, dtype=np.float64)
        _ck(load().mmc_set_chain_state(self.h, chain, _ip(cnt), _dp(com), _dp(off)))

    def _alloc_state(self):
        cap = np.where(self.sys.is_active == 1, self.sys.capacity, self.sys.count)
        return (np.zeros(self.nres, dtype=np.int32), np.zeros((int(cap.sum()) + 1, 3)),
                np.zeros((int((cap * self.sys.natoms).sum()) + 1, 3)))

    def _trim(self, cnt, com, off):
        return cnt, com[:int(cnt.sum())], off[:int((cnt * self.sys.natoms).sum())]

    def get_chain_state(self, chain=0):
        """count, com, offsets in packed order.  Inactive residues are returned from the host copy."""
        cnt, com, off = self._alloc_state()
        _ck(load().mmc_get_chain_state(self.h, chain, _ip(cnt), _dp(com), _dp(off)))
        cnt, com, off = self._trim(cnt, com, off)
        self._fill_inactive(cnt, com, off)
        return cnt, com, off

    def _fill_inactive(self, cnt, com, off):
        mo0, ao0 = self.sys.mol_offsets()
        mo = ao = 0
        for r in range(self.nres):
            n, na = int(cnt[r]), int(self.sys.natoms[r])
            if not self.sys.is_active[r]:
                com[mo:mo + n] = self.sys.com[mo0[r]:mo0[r] + n]
                off[ao:ao + n * na] = self.sys.offsets[ao0[r]:ao0[r] + n * na]
            mo += n; ao += n * na

    def sync_to_host(self, chain=0):
        cnt, com, off = self._alloc_state(); e = np.zeros(MMC_NE)
        _ck(load().mmc_sync_to_host(self.h, chain, _ip(cnt), _dp(com), _dp(off), _dp(e)))
        cnt, com, off = self._trim(cnt, com, off)
        self._fill_inactive(cnt, com, off)
        return cnt, com, off, e

    def reservoir(self, chain=0):
        """(count[nres], offsets [n, 3]) of the chain's reservoir pool"""
        cnt = np.zeros(self.nres, dtype=np.int32)
        _ck(load().mmc_get_reservoir(self.h, chain, _ip(cnt), None))
        n = int((cnt * self.sys.natoms).sum())
        off = np.zeros((max(n, 1), 3))
        _ck(load().mmc_get_reservoir(self.h, chain, _ip(cnt), _dp(off)))
        return cnt, off[:n]

    def sk(self, chain=0):
        a = np.zeros(2 * self.nk)
        _ck(load().mmc_get_sk(self.h, chain, _dp(a)))
        return a[0::2] + 1j * a[1::2]

    # ---- energies
    def system_energy(self, chain=0):
        e = np.zeros(MMC_NE)
        _ck(load().mmc_system_energy(self.h, chain, _dp(e)))
        return e

    def system_energy_all(self):
        """ComputeSystemEnergy of every chain in one launch set (the framework's share is computed once per handle);
        returns [n_chains, 6]"""
        e = np.zeros((self.n_chains, MMC_NE))
        _ck(load().mmc_system_energy_all(self.h, _dp(e)))
        return e

    def trial_energy(self, kind, res, mol, com_new=None, offs_new=None, chain=0):
        old, new = np.zeros(MMC_NE), np.zeros(MMC_NE)
        c = None if com_new is None else np.ascontiguousarray(com_new, dtype=np.float64)
        o = None if offs_new is None else np.ascontiguousarray(offs_new, dtype=np.float64)
        _ck(load().mmc_trial_energy(self.h, chain, kind, res, mol, _dp(c), _dp(o), _dp(old), _dp(new)))
        return old, new

    def trial_energy_swap(self, res_old, mol, res_new, com_new, offs_new, chain=0):
        old, new = np.zeros(MMC_NE), np.zeros(MMC_NE)
        c = np.ascontiguousarray(com_new, dtype=np.float64); o = np.ascontiguousarray(offs_new, dtype=np.float64)
        _ck(load().mmc_trial_energy_swap(self.h, chain, res_old, mol, res_new, _dp(c), _dp(o), _dp(old), _dp(new)))
        return old, new

    def trial_step(self, kind, res, mol, com_new=None, offs_new=None, u=0.0, res_new=-1, chain=0):
        """one strict-mode move in one launch (energies + accept test with the host's uniform u + commit);
        returns (old_e, new_e, prob, accepted)"""
        old, new = np.zeros(MMC_NE), np.zeros(MMC_NE)
        c = None if com_new is None else np.ascontiguousarray(com_new, dtype=np.float64)
        o = None if offs_new is None else np.ascontiguousarray(offs_new, dtype=np.float64)
        p, a = C.c_double(), C.c_int()
        _ck(load().mmc_trial_step(self.h, chain, kind, res, mol, res_new, _dp(c), _dp(o), float(u), _dp(old), _dp(new), C.byref(p), C.byref(a)))
        return old, new, p.value, bool(a.value)

    def trial_commit(self, accept, chain=0):
        _ck(load().mmc_trial_commit(self.h, chain, int(bool(accept))))

    # ---- batched
    def run_batch(self, n_steps, seed=0, tape=None, records=False):
        recs = np.zeros((self.n_chains, n_steps), dtype=RECORD_DTYPE) if records else None
        ctr = (Counters * self.n_chains)()
        stride = 0
        if tape is not None:
            tape = np.ascontiguousarray(tape, dtype=np.float64).reshape(self.n_chains, -1)
            stride = tape.shape[1]
        _ck(load().mmc_run_batch(self.h, n_steps, seed, _dp(tape), stride,
                                 recs.ctypes.data if records else None, C.cast(ctr, C.c_void_p)))
        self.counters = [c.as_dict() for c in ctr]
        return recs

    def run_batch_async(self, n_steps, seed=0):
        _ck(load().mmc_run_batch_async(self.h, n_steps, seed))

    def synchronize(self):
        _ck(load().mmc_synchronize(self.h))

    def tape_used(self, chain=0):
        return load().mmc_tape_used(self.h, chain)

    # ---- bulk per-block exchange (fixed-capacity host arrays, see include/maniac_gpu.h)
    def chain_capacity(self):
        a, b = C.c_int(), C.c_int()
        _ck(load().mmc_chain_capacity(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def upload_chains(self, count, com, offsets):
        """count int32 [C, nres], com f8 [C, g_mols, 3], offsets f8 [C, g_atoms, 3]; numpy arrays or raw
        addresses (ints) of host buffers, e.g. pinned torch tensors' data_ptr()."""
        _ck(load().mmc_upload_chains(self.h, _addr(count), _addr(com), _addr(offsets)))

    def download_chains(self, count=None, com=None, offsets=None, energies=None, counters=None):
        _ck(load().mmc_download_chains(self.h, _addr(count), _addr(com), _addr(offsets), _addr(energies), _addr(counters)))

    def snapshot_begin(self, slot=0):
        """queue a per-block snapshot behind the work already on the stream; returns at once"""
        _ck(load().mmc_snapshot_begin(self.h, slot))

    def snapshot_wait(self, slot=0):
        """(count [C, nres], com [C, g_mols, 3], offsets [C, g_atoms, 3], energies [C, 6], counters [C, N_COUNTERS]) of
        the snapshot on `slot`, copied out of the library's pinned buffers; does not wait for later batches"""
        from .system import N_COUNTERS
        p = [C.c_void_p() for _ in range(5)]
        _ck(load().mmc_snapshot_wait(self.h, slot, *[C.byref(x) for x in p]))
        g_mols, g_atoms = self.chain_capacity()
        n = self.n_chains

        def view(ptr, ctype, shape):
            size = int(np.prod(shape))
            if size == 0:
                return np.zeros(shape, dtype=ctype)
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ctype)), shape=(size,)).reshape(shape).copy()
        return (view(p[0], C.c_int, (n, self.sys.nres)), view(p[1], C.c_double, (n, g_mols, 3)),
                view(p[2], C.c_double, (n, g_atoms, 3)), view(p[3], C.c_double, (n, 6)),
                view(p[4], C.c_int64, (n, N_COUNTERS)))

    def profile_kernels(self, n_steps, seed=0):
        ms = (C.c_double * 3)(); n = (C.c_int64 * 3)()
        _ck(load().mmc_profile_kernels(self.h, n_steps, seed, ms, n))
        return {"pair_ms": ms[0], "recip_ms": ms[1], "tail_ms": ms[2], "launches_each": n[0]}

    # ---- debug
    def dbg_pair_singlemol(self, res, mol, chain=0):
        a, b = C.c_double(), C.c_double()
        _ck(load().mmc_dbg_pair_singlemol(self.h, chain, res, mol, C.byref(a), C.byref(b)))
        return a.value, b.value

    def dbg_self_singlemol(self, res):
        a = C.c_double(); _ck(load().mmc_dbg_self_singlemol(self.h, res, C.byref(a))); return a.value

    def dbg_intra_singlemol(self, res, mol, chain=0):
        a = C.c_double(); _ck(load().mmc_dbg_intra_singlemol(self.h, chain, res, mol, C.byref(a))); return a.value

    def dbg_table_lookup(self, which, r):
        r = np.ascontiguousarray(r, dtype=np.float64); out = np.zeros_like(r)
        _ck(load().mmc_dbg_table_lookup(self.h, which, _dp(r), len(r), _dp(out))); return out

    def dbg_erfc(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64); out = np.zeros_like(x)
        _ck(load().mmc_dbg_erfc(self.h, _dp(x), len(x), _dp(out))); return out

    def erfc_noalloc(self, x):
        """mmc_erfc_noalloc: same results as dbg_erfc without any device allocation (fixed 65536-value scratch per GPU)"""
        x = np.ascontiguousarray(x, dtype=np.float64); out = np.zeros_like(x)
        _ck(load().mmc_erfc_noalloc(self.h, _dp(x), len(x), _dp(out))); return out

    def dbg_set_tracked(self, energies=None, sk=None, chain=0):
        """overwrite tracked energies [6] and / or tracked A(k) (complex [K]) of one chain (test support)"""
        e = None if energies is None else np.ascontiguousarray(energies, dtype=np.float64)
        a = None
        if sk is not None:
            a = np.empty(2 * self.nk); a[0::2] = np.real(sk); a[1::2] = np.imag(sk)
        _ck(load().mmc_dbg_set_tracked(self.h, chain, _dp(e), _dp(a)))

    def stream(self):
        return load().mmc_stream(self.h)

    def chain_launch_config(self):
        """(CTA threads, CTAs per chain) of the batched kernel; CTAs per chain > 1 = the cluster-per-chain kernel"""
        a, b = C.c_int(), C.c_int()
        _ck(load().mmc_chain_launch_config(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def chain_threads(self):
        return self.chain_launch_config()

    def launch_count(self):
        return load().mmc_launch_count(self.h)

    # ---- k-sharded mode (device buffers are raw addresses, e.g. torch tensor.data_ptr())
    def shard_step_begin(self, seed, dev_sums):
        _ck(load().mmc_shard_step_begin(self.h, seed, dev_sums))

    def shard_step_end(self, dev_sums, records=False):
        recs = np.zeros(self.n_chains, dtype=RECORD_DTYPE) if records else None
        _ck(load().mmc_shard_step_end(self.h, dev_sums, recs.ctypes.data if records else None))
        return recs

    def shard_trial_begin(self, kind, res, mol, com_new, offs_new, dev_sums, chain=0):
        c = None if com_new is None else np.ascontiguousarray(com_new, dtype=np.float64)
        o = None if offs_new is None else np.ascontiguousarray(offs_new, dtype=np.float64)
        _ck(load().mmc_shard_trial_begin(self.h, chain, kind, res, mol, _dp(c), _dp(o), dev_sums))

    def shard_trial_end(self, dev_sums, chain=0):
        old, new = np.zeros(MMC_NE), np.zeros(MMC_NE)
        _ck(load().mmc_shard_trial_end(self.h, chain, dev_sums, _dp(old), _dp(new)))
        return old, new

    def shard_range(self):
        a, b = C.c_int(), C.c_int()
        _ck(load().mmc_shard_range(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def strict_loop(self, n_moves, seed=1, fused=True, chain=0):
        """n_moves host-driven translation moves through the strict-mode entry points; (seconds, accepted, launches)"""
        t, a = C.c_double(), C.c_int64()
        l0 = self.launch_count()
        _ck(load().mmc_dbg_strict_loop(self.h, chain, n_moves, seed, int(fused), C.byref(t), C.byref(a)))
        return t.value, int(a.value), self.launch_count() - l0

    def phase_cycles(self):
        """-DMMC_PHASE_TIMING builds: per-phase clock cycles (summed over CTAs and trials) and the trial count"""
        v = (C.c_int64 * 8)()
        _ck(load().mmc_dbg_phase_cycles(self.h, v))
        names = ("propose", "stage", "real_space", "sum4", "reciprocal", "sum2", "finalize_commit", "trials")
        return dict(zip(names, (int(x) for x in v)))

    def fw_cache(self, chain=0):
        """[g_mols, 2] framework-energy cache of one chain by device slot (NaN = nothing cached)"""
        gm, _ = self.chain_capacity()
        out = np.zeros((gm, 2))
        _ck(load().mmc_dbg_fw_cache(self.h, chain, _dp(out)))
        return out

    def pair_work(self):
        """(distances formed, pairs inside reach) by the per-move sweeps since mmc_init"""
        v = (C.c_int64 * 2)()
        _ck(load().mmc_pair_work(self.h, v))
        return int(v[0]), int(v[1])
