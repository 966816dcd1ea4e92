"""Boundary test fed from the reference's REAL array shapes.

The reference keeps its state in column-major module arrays sized for NB_MAX_MOLECULE = 5000 molecules per type
(`input_parser.f90:245-266`, `parameters.f90:8`):
    mol_com(3, nres, NB_MAX)   site_offset(3, nres, NB_MAX, max_atom)   atom_charges/atom_types(nres, max_atom)
    coeff%sigma/epsilon(nres, nres, max_atom, max_atom)
`fortran/maniac_gpu_glue.f90` packs them for `mmc_init` and unpacks `mmc_sync_to_host` back into them; it cannot be
compiled here (no Fortran compiler on either machine), so `tests/fortran_glue_probe.c` transcribes it line for line in C
over numpy arrays in Fortran order.  CPU test: the packed arrays equal the ones the packed path (`System.to_desc`) hands to
the library.  GPU test: mmc_init / mmc_system_energy / mmc_trial_energy / mmc_trial_commit / mmc_sync_to_host driven from
the strided arrays give the packed path's energies and write the state back into the strided arrays.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, load_system, rel_err

NB_MAX_MOLECULE = 5000   # parameters.f90:8
dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int)


class GlueState(C.Structure):
    _fields_ = [("nres", C.c_int), ("max_atom", C.c_int), ("ntypes", C.c_int), ("nb_max", C.c_int),
                ("atom_in_residue", ip), ("is_active", ip), ("num_residues", ip), ("fugacity", dp),
                ("atom_charges", dp), ("atom_types", ip), ("sigma", dp), ("epsilon", dp), ("mol_com", dp), ("site_offset", dp),
                ("bounds_lo", C.c_double * 3), ("matrix_diag", C.c_double * 3), ("tilt", C.c_double * 3),
                ("alpha", C.c_double), ("cutoff", C.c_double), ("temp_K", C.c_double), ("translation_step", C.c_double),
                ("rotation_step_angle", C.c_double), ("p_translation", C.c_double), ("p_rotation", C.c_double), ("p_swap", C.c_double),
                ("kmax", C.c_int * 3), ("energy", C.c_double * 6),
                ("p_natoms", ip), ("p_active", ip), ("p_count", ip), ("p_cap", ip), ("p_types", ip),
                ("p_q", dp), ("p_sig", dp), ("p_eps", dp), ("p_com", dp), ("p_off", dp), ("p_fug", dp),
                ("n_com", C.c_int), ("n_off", C.c_int),
                ("init", C.c_void_p), ("system_energy", C.c_void_p), ("trial_energy", C.c_void_p), ("trial_commit", C.c_void_p),
                ("sync_to_host", C.c_void_p), ("destroy", C.c_void_p), ("gpu", C.c_void_p)]


@pytest.fixture(scope="module")
def probe(tmp_path_factory):
    d = tmp_path_factory.mktemp("glue")
    so = os.path.join(d, "libglueprobe.so")
    subprocess.check_call(["gcc", "-O1", "-shared", "-fPIC", "-I", os.path.join(ROOT, "include"), "-o", so,
                           os.path.join(ROOT, "tests", "fortran_glue_probe.c")])
    L = C.CDLL(so)
    L.glue_sizeof_state.restype = C.c_size_t
    assert L.glue_sizeof_state() == C.sizeof(GlueState)
    for f in ("glue_pack_tables", "glue_pack_coordinates", "glue_detach"):
        getattr(L, f).argtypes = [C.POINTER(GlueState)]; getattr(L, f).restype = None
    L.glue_attach.argtypes = [C.POINTER(GlueState), C.c_int, C.c_int, C.c_int, C.c_int]
    L.glue_system_energy.argtypes = [C.POINTER(GlueState)]
    L.glue_trial.argtypes = [C.POINTER(GlueState), C.c_int, C.c_int, C.c_int, dp, dp]
    L.glue_commit.argtypes = [C.POINTER(GlueState), C.c_int]
    L.glue_sync_primary.argtypes = [C.POINTER(GlueState), C.c_int]
    return L


class FortranState:
    """the reference's module arrays for one System, column-major, allocated at the reference's own extents"""

    def __init__(self, s, nb_max=NB_MAX_MOLECULE):
        self.s = s
        nres, max_atom, ntypes = s.nres, int(s.natoms.max()), s.sigma.shape[0]
        F = dict(order="F")
        self.atom_in_residue = np.asarray(s.natoms, dtype=np.int32).copy()
        self.is_active = np.asarray(s.is_active, dtype=np.int32).copy()
        self.num_residues = np.asarray(s.count, dtype=np.int32).copy()
        self.fugacity = np.asarray(s.fugacity, dtype=np.float64).copy()
        self.atom_charges = np.zeros((nres, max_atom), **F)
        self.atom_types = np.zeros((nres, max_atom), dtype=np.int32, **F)
        for r in range(nres):
            na = int(s.natoms[r])
            self.atom_charges[r, :na] = s.charges[r, :na]; self.atom_types[r, :na] = s.types[r, :na]
        # coeff%sigma/epsilon(res1, res2, atom1, atom2) as parameters_parser.f90 fills them: by the two atoms' types
        T = np.maximum(self.atom_types, 1) - 1
        self.sigma4 = np.asfortranarray(s.sigma[T[:, None, :, None], T[None, :, None, :]])
        self.epsilon4 = np.asfortranarray(s.epsilon[T[:, None, :, None], T[None, :, None, :]])
        assert self.sigma4.shape == (nres, nres, max_atom, max_atom)
        self.mol_com = np.zeros((3, nres, nb_max), **F)
        self.site_offset = np.zeros((3, nres, nb_max, max_atom), **F)
        mo, ao = s.mol_offsets()
        for r in range(nres):
            n, na = int(s.count[r]), int(s.natoms[r])
            self.mol_com[:, r, :n] = s.com[mo[r]:mo[r] + n].T
            self.site_offset[:, r, :n, :na] = s.offsets[ao[r]:ao[r] + n * na].reshape(n, na, 3).transpose(2, 0, 1)
        g = GlueState()
        g.nres, g.max_atom, g.ntypes, g.nb_max = nres, max_atom, ntypes, nb_max
        for name, arr, ct in (("atom_in_residue", self.atom_in_residue, ip), ("is_active", self.is_active, ip),
                              ("num_residues", self.num_residues, ip), ("fugacity", self.fugacity, dp),
                              ("atom_charges", self.atom_charges, dp), ("atom_types", self.atom_types, ip),
                              ("sigma", self.sigma4, dp), ("epsilon", self.epsilon4, dp), ("mol_com", self.mol_com, dp),
                              ("site_offset", self.site_offset, dp)):
            assert arr.flags["F_CONTIGUOUS"]
            setattr(g, name, arr.ctypes.data_as(ct))
        for k in range(3):
            g.bounds_lo[k], g.matrix_diag[k], g.tilt[k], g.kmax[k] = float(s.box_lo[k]), float(s.box_len[k]), float(s.tilt[k]), int(s.kmax[k])
        g.alpha, g.cutoff, g.temp_K = s.alpha, s.cutoff, s.temperature
        g.translation_step, g.rotation_step_angle = s.translation_step, s.rotation_step
        g.p_translation, g.p_rotation, g.p_swap = s.p_translation, s.p_rotation, float(s.p_swap)
        self.g = g

    def bind(self, lib_path):
        L = C.CDLL(lib_path)
        for field, sym in (("init", "mmc_init"), ("system_energy", "mmc_system_energy"), ("trial_energy", "mmc_trial_energy"),
                           ("trial_commit", "mmc_trial_commit"), ("sync_to_host", "mmc_sync_to_host"), ("destroy", "mmc_destroy")):
            setattr(self.g, field, C.cast(getattr(L, sym), C.c_void_p).value)
        self._lib = L


def view(ptr, n, dtype):
    ct = C.c_int if dtype == np.int32 else C.c_double
    return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct)), shape=(max(n, 1),))[:n].copy()


@pytest.mark.parametrize("name", ["lj_gas", "h2o_ads", "slit"])
def test_glue_packing_equals_packed_path(probe, name):
    """pack_coordinates + the packing half of gpu_attach over the strided Fortran arrays == System.to_desc's arrays"""
    s = load_system(name)
    f = FortranState(s)
    probe.glue_pack_tables(C.byref(f.g))
    g, d = f.g, s.to_desc()
    nres, maxat, nt = s.nres, int(s.natoms.max()), s.sigma.shape[0]
    assert np.array_equal(view(g.p_natoms, nres, np.int32), s.natoms) and np.array_equal(view(g.p_active, nres, np.int32), s.is_active)
    assert np.array_equal(view(g.p_count, nres, np.int32), s.count)
    cap = view(g.p_cap, nres, np.int32)
    assert np.all(cap[s.is_active == 1] == NB_MAX_MOLECULE) and np.array_equal(cap[s.is_active != 1], s.count[s.is_active != 1])
    q, ty = view(g.p_q, nres * maxat, np.float64).reshape(nres, maxat), view(g.p_types, nres * maxat, np.int32).reshape(nres, maxat)
    for r in range(nres):
        na = int(s.natoms[r])
        assert np.array_equal(q[r, :na], s.charges[r, :na]) and np.array_equal(ty[r, :na], s.types[r, :na])
    used = np.zeros((nt, nt), dtype=bool)      # the glue fills every type pair that occurs between two residues' atoms
    tt = [int(s.types[r, a]) - 1 for r in range(nres) for a in range(int(s.natoms[r]))]
    used[np.ix_(tt, tt)] = True
    sig, eps = view(g.p_sig, nt * nt, np.float64).reshape(nt, nt), view(g.p_eps, nt * nt, np.float64).reshape(nt, nt)
    assert np.array_equal(sig[used], s.sigma[used]) and np.array_equal(eps[used], s.epsilon[used])
    assert g.n_com == int(s.count.sum()) and g.n_off == int((s.count * s.natoms).sum())
    assert np.array_equal(view(g.p_com, 3 * g.n_com, np.float64), d._keep["com"].reshape(-1)[:3 * g.n_com])
    assert np.array_equal(view(g.p_off, 3 * g.n_off, np.float64), d._keep["offsets"].reshape(-1)[:3 * g.n_off])
    probe.glue_detach(C.byref(f.g))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["zif8_h2o", "lj_gas"])
def test_abi_driven_from_fortran_layouts(probe, name):
    """The reference's own C3 example at its real extents (coeff%sigma/epsilon(2,2,2208,2208) = 156 MB each,
    site_offset(3,2,5000,2208)): gpu_attach -> gpu_system_energy -> gpu_trial / gpu_commit -> gpu_sync_primary through the
    C transcription of the Fortran glue == the packed path (ManiacGPU) on the same moves, and the write-back lands in the
    strided arrays."""
    from maniac_mc_b200 import lib
    from maniac_mc_b200.lib import ManiacGPU
    from maniac_mc_b200.system import MMC_CREATION, MMC_DELETION, MMC_TRANSLATION
    s = load_system(name)                      # reference-default quirk switches, as the glue passes them
    s.capacity = np.where(s.is_active == 1, NB_MAX_MOLECULE, s.count).astype(np.int32)
    f = FortranState(s); f.bind(lib.SO_PATH)
    g = f.g
    assert probe.glue_attach(C.byref(g), 1, 0, 0, 0) == 0, lib.load().mmc_last_error()
    ref = ManiacGPU(s)
    assert probe.glue_system_energy(C.byref(g)) == 0
    e_ref = ref.system_energy()
    assert rel_err(np.array(g.energy[:]), e_ref) < 1e-12
    r = int(np.flatnonzero(s.is_active == 1)[0]); na = int(s.natoms[r]); n0 = int(s.count[r])
    mo, ao = s.mol_offsets()
    old, new = (C.c_double * 6)(), (C.c_double * 6)()
    # translation of molecule 1 of the active type: the Fortran move routine writes the proposal into mol_com first
    com_old = f.mol_com[:, r, 0].copy()
    f.mol_com[:, r, 0] = com_old + np.array([0.31, -0.22, 0.17])
    assert probe.glue_trial(C.byref(g), MMC_TRANSLATION, r + 1, 1, old, new) == 0, lib.load().mmc_last_error()
    o2, n2 = ref.trial_energy(MMC_TRANSLATION, r, 0, f.mol_com[:, r, 0].copy(), s.offsets[ao[r]:ao[r] + na])
    assert rel_err(np.array(old[:]), o2) < 1e-12 and rel_err(np.array(new[:]), n2) < 1e-12
    assert probe.glue_commit(C.byref(g), 1) == 0; ref.trial_commit(True)
    # creation in slot count+1 (geometry written into the arrays first, create_molecule.f90:180-199), then accept
    f.mol_com[:, r, n0] = s.box_lo + 0.37 * s.box_len
    f.site_offset[:, r, n0, :na] = f.site_offset[:, r, 0, :na]
    assert probe.glue_trial(C.byref(g), MMC_CREATION, r + 1, n0 + 1, old, new) == 0, lib.load().mmc_last_error()
    o2, n2 = ref.trial_energy(MMC_CREATION, r, n0, f.mol_com[:, r, n0].copy(), np.ascontiguousarray(f.site_offset[:, r, n0, :na].T))
    assert rel_err(np.array(old[:]), o2) < 1e-12 and rel_err(np.array(new[:]), n2) < 1e-12
    assert probe.glue_commit(C.byref(g), 1) == 0; ref.trial_commit(True)
    # a rejected deletion changes nothing
    assert probe.glue_trial(C.byref(g), MMC_DELETION, r + 1, 1, old, new) == 0
    assert probe.glue_commit(C.byref(g), 0) == 0
    # write-back: scribble over the host arrays, then gpu_sync_primary must restore the device's state into them
    want_com = f.mol_com[:, r, :n0 + 1].copy(); want_off = f.site_offset[:, r, :n0 + 1, :na].copy()
    f.mol_com[:, r, :n0 + 1] = -999.0; f.site_offset[:, r, :n0 + 1, :na] = -999.0; f.num_residues[r] = -1
    fw = [q for q in range(s.nres) if s.is_active[q] != 1]
    fw_before = [f.mol_com[:, q, :int(s.count[q])].copy() for q in fw]
    assert probe.glue_sync_primary(C.byref(g), 0) == 0, lib.load().mmc_last_error()
    assert f.num_residues[r] == n0 + 1
    assert np.array_equal(f.mol_com[:, r, :n0 + 1], want_com) and np.array_equal(f.site_offset[:, r, :n0 + 1, :na], want_off)
    for q, b in zip(fw, fw_before):
        assert np.array_equal(f.mol_com[:, q, :int(s.count[q])], b)          # inactive residues are never written
    cnt, com, off, e = ref.sync_to_host()
    assert rel_err(np.array(g.energy[:]), e) < 1e-12 and cnt[r] == n0 + 1
    probe.glue_detach(C.byref(g)); ref.close()
