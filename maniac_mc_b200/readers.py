"""Minimal readers for MANIAC's three input files (.maniac, LAMMPS .data, pair_coeff include).

In the drop-in deployment these files are read by the reference's own Fortran parsers
(input_parser.f90, data_parser.f90, parameters_parser.f90), which stay untouched; the Fortran
shim then packs ``simulation_state`` into ``mmc_desc``.  This module exists for the host-side
harness in this repository (tests, bench, fixture generation): it produces the same packed
description from the same files.  It is host set-up code, not part of the device hot path.

Followed reference behaviour (paths relative to the reference's src/):
  * keyword file            input_parser.f90:297-601 (unknown keywords ignored, Q14)
  * probability rescale      input_parser.f90:73-100
  * residues sorted by their smallest atom type   input_parser.f90:603-672
  * atoms sorted by id       data_parser.f90:1060-1125
  * molecule detection by type pattern            data_parser.f90:1205-1287
  * active molecules made whole across PBC        data_parser.f90:1290-1385, readers_utils.f90:258-310
  * COM (uniform weights, see note) + offsets     data_parser.f90:1386-1511
  * pair_coeff table, epsilon kcal/mol -> K       parameters_parser.f90:60-103
  * Lorentz-Berthelot fill-in of empty pairs      parameters_parser.f90:116-182
  * fugacity atm -> 1/A^3    prepare_utils.f90:48-75
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

KB_JK = 1.380658e-23          # constants.f90:10
KB_KCALMOL = 0.0019872041     # constants.f90:11
ATM_TO_PA = 1.01325e5         # parameters.f90:29
A3_TO_M3 = 1.0e-30            # parameters.f90:28
NB_MAX_MOLECULE = 5000        # parameters.f90:8


@dataclass
class Residue:
    name: str
    active: bool
    fugacity_atm: float
    types: list
    natoms: int


@dataclass
class ManiacInput:
    nb_block: int = 0
    nb_step: int = 0
    temperature: float = 300.0
    ewald_tolerance: float = 1e-5
    real_space_cutoff: float = 10.0
    translation_step: float = 1.0
    rotation_step_angle: float = 0.5
    recalibrate_moves: bool = False
    p_translation: float = 0.0
    p_rotation: float = 0.0
    p_insertion_deletion: float = 0.0
    p_swap: float = 0.0
    seed: int = 0
    residues: list = field(default_factory=list)


def read_maniac_input(path) -> ManiacInput:
    inp = ManiacInput()
    cur = None
    with open(path) as fh:
        for raw in fh:
            line = raw.split("#", 1)[0].strip()
            if not line:
                continue
            tok = line.split()
            key, vals = tok[0], tok[1:]
            if key == "begin_residue":
                cur = dict(name="", active=False, fugacity=0.0, types=[], natoms=0)
                continue
            if key == "end_residue":
                inp.residues.append(Residue(cur["name"], cur["active"], cur["fugacity"], cur["types"], cur["natoms"]))
                cur = None
                continue
            if cur is not None:
                if key == "name":
                    cur["name"] = vals[0]
                elif key == "state":
                    cur["active"] = vals[0].lower().startswith("act")
                elif key == "fugacity":
                    cur["fugacity"] = float(vals[0])
                elif key == "types":
                    cur["types"] = [int(v) for v in vals]
                elif key == "nb-atoms":
                    cur["natoms"] = int(vals[0])
                continue
            if key == "nb_block":
                inp.nb_block = int(vals[0])
            elif key == "nb_step":
                inp.nb_step = int(vals[0])
            elif key == "temperature":
                inp.temperature = float(vals[0])
            elif key == "ewald_tolerance":
                inp.ewald_tolerance = float(vals[0])
            elif key == "real_space_cutoff":
                inp.real_space_cutoff = float(vals[0])
            elif key == "translation_step":
                inp.translation_step = float(vals[0])
            elif key == "rotation_step_angle":
                inp.rotation_step_angle = float(vals[0])
            elif key == "recalibrate_moves":
                inp.recalibrate_moves = vals[0].lower() in ("true", ".true.", "1", "yes")
            elif key == "translation_proba":
                inp.p_translation = float(vals[0])
            elif key == "rotation_proba":
                inp.p_rotation = float(vals[0])
            elif key == "insertion_deletion_proba":
                inp.p_insertion_deletion = float(vals[0])
            elif key == "swap_proba":
                inp.p_swap = float(vals[0])
            elif key == "seed":
                inp.seed = int(vals[0])
            # anything else (e.g. big_move_proba) is silently ignored like the reference does
    tot = inp.p_translation + inp.p_rotation + inp.p_insertion_deletion + inp.p_swap
    if tot > 0 and abs(tot - 1.0) > 1e-10:
        s = 1.0 / tot
        inp.p_translation *= s
        inp.p_rotation *= s
        inp.p_insertion_deletion *= s
        inp.p_swap *= s
    inp.residues.sort(key=lambda r: min(r.types))  # stable, like the insertion sort in SortResidues
    return inp


def read_lammps_data(path):
    """Returns dict(box_lo, box_hi, tilt (xy, xz, yz), masses{type:mass}, atoms=array[id, mol, type, q, x, y, z])."""
    lo, hi = [0.0] * 3, [0.0] * 3
    tilt = [0.0, 0.0, 0.0]
    masses, atoms = {}, []
    natoms = ntypes = 0
    section = None
    with open(path) as fh:
        lines = fh.readlines()
    for raw in lines[1:]:
        line = raw.split("#", 1)[0].strip()
        if not line:
            continue
        tok = line.split()
        if line.endswith("atoms") and len(tok) == 2:
            natoms = int(tok[0]); continue
        if line.endswith("atom types"):
            ntypes = int(tok[0]); continue
        if len(tok) == 4 and tok[2] in ("xlo", "ylo", "zlo"):
            d = "xyz".index(tok[2][0]); lo[d], hi[d] = float(tok[0]), float(tok[1]); continue
        if len(tok) >= 6 and tok[3] == "xy":      # "xy xz yz" tilt factors (ParseLAMMPSBox readers_utils.f90:205-214)
            tilt = [float(v) for v in tok[:3]]
            continue
        if tok[0][0].isalpha():
            section = " ".join(tok); continue
        if section == "Masses":
            masses[int(tok[0])] = float(tok[1])
        elif section == "Atoms":
            # atom_style full: id mol type q x y z [ix iy iz]
            atoms.append((int(tok[0]), int(tok[1]), int(tok[2]), float(tok[3]), float(tok[4]), float(tok[5]), float(tok[6])))
    if not all(h > l for l, h in zip(lo, hi)):
        raise ValueError(f"{path}: missing or empty box bounds")
    if natoms == 0 or len(atoms) != natoms:
        raise ValueError(f"{path}: expected {natoms} atoms, found {len(atoms)}")
    if not masses or len(masses) < ntypes:
        raise ValueError(f"{path}: missing or incomplete Masses section")
    a = np.array(sorted(atoms, key=lambda t: t[0]), dtype=np.float64)
    return dict(box_lo=np.array(lo), box_hi=np.array(hi), tilt=np.array(tilt), masses=masses, atoms=a, ntypes=ntypes)


def read_pair_coeffs(path, ntypes):
    """sigma[ti,tj] (A), epsilon[ti,tj] (K); unspecified pairs are zero before the LB fill-in."""
    sigma = np.zeros((ntypes, ntypes)); eps = np.zeros((ntypes, ntypes))
    with open(path) as fh:
        for raw in fh:
            line = raw.split("#", 1)[0].strip()
            if not line.startswith("pair_coeff"):
                continue
            tok = line.split()
            i, j, e, s = int(tok[1]) - 1, int(tok[2]) - 1, float(tok[3]), float(tok[4])
            if i >= ntypes or j >= ntypes:
                continue
            e = e / KB_KCALMOL
            sigma[i, j] = sigma[j, i] = s
            eps[i, j] = eps[j, i] = e
    return sigma, eps


def lorentz_berthelot(sigma, eps):
    """ApplyLorentzBerthelot, parameters_parser.f90:116-182: only pairs with both entries ~0."""
    n = sigma.shape[0]
    s2, e2 = sigma.copy(), eps.copy()
    for i in range(n):
        for j in range(n):
            if abs(eps[i, j]) < 1e-6 and abs(sigma[i, j]) < 1e-6:
                s = (sigma[i, i] + sigma[j, j]) / 2
                e = math.sqrt(eps[i, i] * eps[j, j])
                if s > 1e-6 and e > 1e-6:
                    s2[i, j], e2[i, j] = s, e
    return s2, e2


def wrap_nearest(x, L):
    return np.mod(x + 0.5 * L, L) - 0.5 * L


def nearest_image_tri(d, L, tilt):
    """Nearest lattice image of separations d [..., 3] in the cell a = (lx,0,0), b = (xy,ly,0), c = (xz,yz,lz):
    cascade z -> y -> x into the brick, then the 27 neighbouring images (same definition as the library's)."""
    d = np.array(d, dtype=np.float64)
    xy, xz, yz = tilt
    n = np.rint(d[..., 2] / L[2]); d[..., 2] -= n * L[2]; d[..., 1] -= n * yz; d[..., 0] -= n * xz
    n = np.rint(d[..., 1] / L[1]); d[..., 1] -= n * L[1]; d[..., 0] -= n * xy
    n = np.rint(d[..., 0] / L[0]); d[..., 0] -= n * L[0]
    best = d.copy(); best2 = (d ** 2).sum(-1)
    for sa in (-1, 0, 1):
        for sb in (-1, 0, 1):
            for sc in (-1, 0, 1):
                t = d + np.array([sa * L[0] + sb * xy + sc * xz, sb * L[1] + sc * yz, sc * L[2]])
                r2 = (t ** 2).sum(-1)
                m = r2 < best2
                best[m] = t[m]; best2[m] = r2[m]
    return best


def wrap_tri(p, lo, L, tilt):
    """ApplyPBC for a tilted cell: s = (a b c)^-1 (p - lo), s mod 1, back to Cartesian"""
    xy, xz, yz = tilt
    r = np.asarray(p, dtype=np.float64) - lo
    sz = r[..., 2] / L[2]; sy = (r[..., 1] - yz * sz) / L[1]; sx = (r[..., 0] - xy * sy - xz * sz) / L[0]
    sx, sy, sz = sx - np.floor(sx), sy - np.floor(sy), sz - np.floor(sz)
    return lo + np.stack([L[0] * sx + xy * sy + xz * sz, L[1] * sy + yz * sz, L[2] * sz], axis=-1)
