"""Builders of ``System``: from the reference's input files, and synthetic benchmark systems
(C2 TIP4P box, C4 ~50k-atom framework + guests)."""
from __future__ import annotations

import math

import numpy as np

from . import readers
from .system import System, ewald_setup


def from_files(maniac_path, data_path, inc_path, capacity=None) -> System:
    """ReadInput + ReadSystemData + ReadParameters + PrepareSimulationParameters (main.f90:17-25)
    reduced to what the energy path needs."""
    inp = readers.read_maniac_input(maniac_path)
    dat = readers.read_lammps_data(data_path)
    atoms = dat["atoms"]
    ntypes = max(dat["ntypes"], int(atoms[:, 2].max()))
    lo, hi = dat["box_lo"], dat["box_hi"]
    L = hi - lo
    tilt = np.asarray(dat.get("tilt", np.zeros(3)), dtype=np.float64)
    tri = bool(np.any(tilt != 0.0))
    nres = len(inp.residues)
    max_atoms = max(r.natoms for r in inp.residues)
    natoms = np.array([r.natoms for r in inp.residues], dtype=np.int32)
    is_active = np.array([1 if r.active else 0 for r in inp.residues], dtype=np.int32)
    count = np.zeros(nres, dtype=np.int32)
    charges = np.zeros((nres, max_atoms)); types = np.zeros((nres, max_atoms), dtype=np.int32)
    coms, offs = [], []
    atype = atoms[:, 2].astype(int)
    for ri, res in enumerate(inp.residues):
        sel = np.isin(atype, res.types)
        a = atoms[sel]
        if len(a) % res.natoms != 0:
            raise ValueError(f"residue {res.name}: {len(a)} atoms is not a multiple of nb-atoms {res.natoms}")
        n = len(a) // res.natoms
        count[ri] = n
        if n == 0:
            # no molecule of this type in the box: template must come from a reservoir (Q9)
            continue
        xyz = a[:, 4:7].reshape(n, res.natoms, 3).copy()
        charges[ri, :res.natoms] = a[:res.natoms, 3]
        types[ri, :res.natoms] = a[:res.natoms, 2].astype(int)
        if res.active:
            # RepairMolecule readers_utils.f90:258: chain of nearest-image bonds
            for k in range(1, res.natoms):
                dk = xyz[:, k] - xyz[:, k - 1]
                xyz[:, k] = xyz[:, k - 1] + (readers.nearest_image_tri(dk, L, tilt) if tri else readers.wrap_nearest(dk, L))
        # TransformCoordinate data_parser.f90:1386: the mass vector is overwritten by a scalar
        # (:1421), so the "COM" is the unweighted centroid
        com0 = xyz.mean(axis=1) if res.natoms > 1 else xyz[:, 0].copy()
        if res.natoms > 1:
            com0 = np.zeros((n, 3))
            for k in range(res.natoms):
                com0 = com0 + 1.0 * xyz[:, k]
            com0 = com0 / float(res.natoms)
        off = xyz - com0[:, None, :]
        com = readers.wrap_tri(com0, lo, L, tilt) if tri else lo + np.mod(com0 - lo, L)  # ApplyPBC geometry_utils.f90:167
        coms.append(com); offs.append(off.reshape(-1, 3))
    sigma, eps = readers.read_pair_coeffs(inc_path, ntypes)
    sigma, eps = readers.lorentz_berthelot(sigma, eps)
    cutoff = inp.real_space_cutoff
    # box%metrics(1:3), geometry_utils.f90:124: column norms of the matrix whose rows are a, b, c (= L when not tilted)
    metrics = np.array([math.sqrt(L[0] ** 2 + tilt[0] ** 2 + tilt[1] ** 2), math.sqrt(L[1] ** 2 + tilt[2] ** 2), L[2]]) if tri else L
    if (cutoff > metrics).any():  # AdjustRealSpaceCutoff prepare_utils.f90:141-150
        cutoff = float(metrics.min()) / 2.0
    alpha, kmax = ewald_setup(cutoff, inp.ewald_tolerance, metrics)
    # ConvertFugacity prepare_utils.f90:48-75
    fug = np.array([(r.fugacity_atm * readers.ATM_TO_PA * readers.A3_TO_M3 / (readers.KB_JK * inp.temperature))
                    if r.active else r.fugacity_atm for r in inp.residues])
    if capacity is None:
        capacity = np.where(is_active == 1, readers.NB_MAX_MOLECULE, count).astype(np.int32)
    else:
        capacity = np.where(is_active == 1, np.maximum(capacity, count), count).astype(np.int32)
    return System(natoms=natoms, is_active=is_active, count=count, capacity=capacity, charges=charges, types=types,
                  sigma=sigma, epsilon=eps, com=np.concatenate(coms) if coms else np.zeros((0, 3)),
                  offsets=np.concatenate(offs) if offs else np.zeros((0, 3)),
                  box_lo=lo.copy(), box_len=L.copy(), alpha=alpha, cutoff=cutoff, temperature=inp.temperature,
                  fugacity=fug, kmax=kmax, translation_step=inp.translation_step,
                  rotation_step=inp.rotation_step_angle, p_translation=inp.p_translation,
                  p_rotation=inp.p_rotation, p_swap=inp.p_swap, names=[r.name for r in inp.residues],
                  ewald_tolerance=inp.ewald_tolerance, tilt=tilt)


# ---- TIP4P/2005-like water of mc-topology/molecule-reservoir/tip4p-water (values restated) ----
TIP4P_Q = np.array([0.0, -1.1128, 0.5564, 0.5564])     # O, M, H, H
TIP4P_EPS_OO_KCAL = 0.155062
TIP4P_SIG_OO = 3.154


def tip4p_geometry():
    """O, M, H, H offsets from the unweighted centroid; r_OH=0.9572, HOH=104.52 deg, r_OM=0.15 (bond_coeff/angle_coeff
    of the reference's water.inc)."""
    roh, ang, rom = 0.9572, math.radians(104.52), 0.15
    O = np.zeros(3)
    H1 = np.array([roh * math.sin(ang / 2), roh * math.cos(ang / 2), 0.0])
    H2 = np.array([-roh * math.sin(ang / 2), roh * math.cos(ang / 2), 0.0])
    M = np.array([0.0, rom, 0.0])
    xyz = np.stack([O, M, H1, H2])
    return xyz - xyz.mean(axis=0)


def random_rotations(rng, n):
    q = rng.normal(size=(n, 4)); q /= np.linalg.norm(q, axis=1)[:, None]
    a, b, c, d = q.T
    return np.stack([np.stack([a * a + b * b - c * c - d * d, 2 * (b * c - a * d), 2 * (b * d + a * c)], -1),
                     np.stack([2 * (b * c + a * d), a * a - b * b + c * c - d * d, 2 * (c * d - a * b)], -1),
                     np.stack([2 * (b * d - a * c), 2 * (c * d + a * b), a * a - b * b - c * c + d * d], -1)], 1)


def fugacity_from_atm(f_atm, T):
    return f_atm * readers.ATM_TO_PA * readers.A3_TO_M3 / (readers.KB_JK * T)


def synthetic_framework(n_framework=50000, n_guests=200, box=96.0, cutoff=17.0, tol=1e-5, seed=4,
                        capacity=512, temperature=300.0, fugacity_atm=1.0e4, min_sep=1.5,
                        p_translation=0.4, p_rotation=0.4) -> System:
    """C4: n_framework atoms on a jittered lattice in a cubic box (>= min_sep apart), 8 framework atom
    types cycled with ZIF-8-like LJ parameters and charges shifted to zero net charge, as ONE inactive residue
    (Q13), plus n_guests rigid TIP4P molecules placed in lattice voids."""
    rng = np.random.default_rng(seed)
    # framework types 4..11 of zif8-water.inc (epsilon kcal/mol, sigma A), restated values
    fw_eps = np.array([0.085617, 0.072499, 0.072499, 0.030380, 0.072499, 0.030380, 0.030380, 0.047641])
    fw_sig = np.array([2.338, 3.259, 3.259, 2.440, 3.259, 2.440, 2.440, 3.097])
    fw_q = np.array([0.6918, 0.4291, -0.0839, 0.1128, -0.4526, 0.1325, 0.1306, -0.3879])
    g = int(math.ceil((n_framework + 2 * n_guests) ** (1 / 3)))
    a = box / g
    idx = rng.permutation(g ** 3)
    cells = np.stack(np.unravel_index(idx, (g, g, g)), -1).astype(float)
    jitter = max(0.0, (a - min_sep) / 2.0)
    fw_pos = (cells[:n_framework] + 0.5) * a + rng.uniform(-jitter, jitter, size=(n_framework, 3)) * 0.9
    fw_types = (np.arange(n_framework) % 8)
    q = fw_q[fw_types].copy(); q -= q.mean()
    ntypes = 3 + 8
    eps_d = np.concatenate([[TIP4P_EPS_OO_KCAL, 0.0, 0.0], fw_eps]) / readers.KB_KCALMOL
    sig_d = np.concatenate([[TIP4P_SIG_OO, 0.0, 0.0], fw_sig])
    sigma = np.zeros((ntypes, ntypes)); eps = np.zeros((ntypes, ntypes))
    for i in range(ntypes):
        for j in range(ntypes):
            if eps_d[i] > 0 and eps_d[j] > 0:
                sigma[i, j] = 0.5 * (sig_d[i] + sig_d[j]); eps[i, j] = math.sqrt(eps_d[i] * eps_d[j])
    geo = tip4p_geometry()
    R = random_rotations(rng, n_guests)
    g_com = (cells[n_framework:n_framework + n_guests] + 0.5) * a
    g_off = np.einsum("nij,aj->nai", R, geo)
    max_atoms = n_framework
    charges = np.zeros((2, max_atoms)); types = np.zeros((2, max_atoms), dtype=np.int32)
    charges[0, :4] = TIP4P_Q; types[0, :4] = [1, 2, 3, 3]
    charges[1, :] = q; types[1, :] = fw_types + 4
    fw_com = fw_pos.mean(axis=0)
    L = np.array([box] * 3); lo = np.zeros(3)
    alpha, kmax = ewald_setup(cutoff, tol, L)
    return System(natoms=np.array([4, n_framework], dtype=np.int32), is_active=np.array([1, 0], dtype=np.int32),
                  count=np.array([n_guests, 1], dtype=np.int32), capacity=np.array([max(capacity, n_guests), 1], dtype=np.int32),
                  charges=charges, types=types, sigma=sigma, epsilon=eps,
                  com=np.concatenate([g_com, fw_com[None]]), offsets=np.concatenate([g_off.reshape(-1, 3), fw_pos - fw_com]),
                  box_lo=lo, box_len=L, alpha=alpha, cutoff=cutoff, temperature=temperature,
                  fugacity=np.array([fugacity_from_atm(fugacity_atm, temperature), 0.0]), kmax=kmax,
                  translation_step=0.5, rotation_step=0.5, p_translation=p_translation, p_rotation=p_rotation,
                  names=["wat", "fw"], ewald_tolerance=tol)


def synthetic_water_box(n_mol=64, box=30.0, cutoff=14.0, tol=1e-5, seed=1, capacity=1024, temperature=300.0,
                        fugacity_atm=0.05, p_translation=0.25, p_rotation=0.25) -> System:
    """C2: rigid TIP4P molecules on a lattice in an otherwise empty cubic box."""
    rng = np.random.default_rng(seed)
    g = int(math.ceil(n_mol ** (1 / 3)))
    a = box / g
    cells = np.stack(np.unravel_index(rng.permutation(g ** 3)[:n_mol], (g, g, g)), -1).astype(float)
    com = (cells + 0.5) * a + rng.uniform(-0.2, 0.2, size=(n_mol, 3))
    off = np.einsum("nij,aj->nai", random_rotations(rng, n_mol), tip4p_geometry())
    sigma = np.zeros((3, 3)); eps = np.zeros((3, 3))
    sigma[0, 0] = TIP4P_SIG_OO; eps[0, 0] = TIP4P_EPS_OO_KCAL / readers.KB_KCALMOL
    L = np.array([box] * 3)
    alpha, kmax = ewald_setup(cutoff, tol, L)
    return System(natoms=np.array([4], dtype=np.int32), is_active=np.array([1], dtype=np.int32),
                  count=np.array([n_mol], dtype=np.int32), capacity=np.array([max(capacity, n_mol)], dtype=np.int32),
                  charges=TIP4P_Q[None].copy(), types=np.array([[1, 2, 3, 3]], dtype=np.int32), sigma=sigma, epsilon=eps,
                  com=com, offsets=off.reshape(-1, 3), box_lo=np.zeros(3), box_len=L, alpha=alpha, cutoff=cutoff,
                  temperature=temperature, fugacity=np.array([fugacity_from_atm(fugacity_atm, temperature)]), kmax=kmax,
                  p_translation=p_translation, p_rotation=p_rotation, names=["wat"], ewald_tolerance=tol)
