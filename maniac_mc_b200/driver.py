"""Strict-mode host driver: the reference's move routines, calling the device through the C ABI.

Mirrors, name for name, the Fortran routines that stay on the host in the drop-in deployment
(they keep drawing the random numbers and making the proposals; only the energy work moves to
the GPU):

    MonteCarloLoop            monte_carlo.f90:26-88
    Translation               translation.f90:36-112
    Rotation                  rotation.f90:34-75
    CreateMolecule            create_molecule.f90:41-207
    DeleteMolecule            delete_molecule.f90:41-199
    PickRandomResidueType / PickRandomMoleculeIndex / ApplyRandomRotation / AdjustMoveStepSizes /
    mc_acceptance_probability monte_carlo_utils.f90:30-221

Random numbers come from a callable ``rng()`` returning uniforms in [0,1) (a recorded tape in
the tests; the reference's ``random_number`` stream is not reproducible).
"""
from __future__ import annotations

import math

import numpy as np

from .lib import ManiacGPU
from .system import (E_COUL, E_INTRA, E_NONCOUL, E_RECIP, E_SELF, E_TOTAL, MMC_CREATION, MMC_DELETION, MMC_ROTATION, MMC_SWAP,
                     MMC_TRANSLATION, RECORD_DTYPE, System)

TWOPI = 2.0 * 3.14159265358979323846
PROB_CREATE_DELETE = 0.5
TARGET_ACCEPTANCE, TOL_ACCEPTANCE = 0.40, 0.05
MIN_TRANSLATION_STEP, MAX_TRANSLATION_STEP = 1.0e-3, 3.0
MIN_ROTATION_ANGLE, MAX_ROTATION_ANGLE = 1.0e-3, 0.78
MIN_TRIALS_FOR_RECALIBRATION = 500


class TapeRng:
    def __init__(self, tape):
        self.tape = np.asarray(tape, dtype=np.float64); self.pos = 0

    def __call__(self):
        v = self.tape[self.pos]; self.pos += 1
        return float(v)


def f_modulo(a, p):
    r = math.fmod(a, p)
    if r != 0.0 and ((r < 0.0) != (p < 0.0)):
        r += p
    return r


class StrictDriver:
    """Holds the host mirror of primary%mol_com / site_offset / num_residues for the ACTIVE residues
    exactly like the Fortran driver does, and the ``energy`` / ``counter`` globals."""

    def __init__(self, system: System, gpu: ManiacGPU | None = None, chain=0, recalibrate=False, fused=False):
        self.sys = system
        self.gpu = gpu if gpu is not None else ManiacGPU(system)
        self.chain = chain
        s = system
        mo, ao = s.mol_offsets()
        self.count = s.count.copy()
        self.com, self.off = [], []
        for r in range(s.nres):
            cap, na, n = int(s.capacity[r]), int(s.natoms[r]), int(s.count[r])
            com = np.zeros((max(cap, n), 3)); off = np.zeros((max(cap, n), na, 3))
            com[:n] = s.com[mo[r]:mo[r] + n]; off[:n] = s.offsets[ao[r]:ao[r] + n * na].reshape(n, na, 3)
            self.com.append(com); self.off.append(off)
        self.translation_step, self.rotation_step = s.translation_step, s.rotation_step
        self.recalibrate = recalibrate
        # fused: one mmc_trial_step per move (energies + accept test with the uniform drawn here + commit in ONE launch) instead of
        # mmc_trial_energy + host decision + mmc_trial_commit; the uniform is drawn at the same place of the stream either way
        self.fused = fused
        self.counter = dict(translations=0, rotations=0, creations=0, deletions=0, trial_translations=0,
                            trial_rotations=0, trial_creations=0, trial_deletions=0, swaps=0, trial_swaps=0)
        self.energy = np.zeros(6)
        # reservoir box (has_reservoir): in strict mode the host owns the pool, as the Fortran driver does
        self.has_reservoir = s.reservoir_count is not None
        self.rsv = []
        if self.has_reservoir:
            ro = np.asarray(s.reservoir_offsets, dtype=np.float64).reshape(-1, 3); a0 = 0
            for r in range(s.nres):
                n, na = int(s.reservoir_count[r]), int(s.natoms[r])
                self.rsv.append([ro[a0 + i * na:a0 + (i + 1) * na].copy() for i in range(n)]); a0 += n * na

    # ComputeSystemEnergy energy_utils.f90:18
    def ComputeSystemEnergy(self):
        self.energy = self.gpu.system_energy(self.chain)
        return self.energy

    # mc_acceptance_probability monte_carlo_utils.f90:184-221
    def mc_acceptance_probability(self, old, new, r, move):
        N = float(self.count[r]); V = self.sys.volume; phi = float(self.sys.fugacity[r]); T = self.sys.temperature
        de = new[E_TOTAL] - old[E_TOTAL]
        try:
            ex = math.exp(-de / T)
        except OverflowError:
            ex = math.inf
        if move == MMC_CREATION:
            return min(1.0, (phi * V / N) * ex)
        if move == MMC_DELETION:
            return min(1.0, ((N + 1.0) / (phi * V)) * ex)
        return min(1.0, ex)

    def _accept_move(self, old, new):  # AcceptMove :410
        e = self.energy
        e[E_RECIP] = e[E_RECIP] + new[E_RECIP] - old[E_RECIP]
        e[E_NONCOUL] = e[E_NONCOUL] + new[E_NONCOUL] - old[E_NONCOUL]
        e[E_COUL] = e[E_COUL] + new[E_COUL] - old[E_COUL]
        e[E_TOTAL] = e[E_TOTAL] + new[E_TOTAL] - old[E_TOTAL]

    def _accept_insdel(self, old, new):  # AcceptCreationMove / AcceptDeletionMove
        e = self.energy
        e[E_RECIP] = new[E_RECIP]
        for k in (E_NONCOUL, E_COUL, E_SELF, E_INTRA, E_TOTAL):
            e[k] = e[k] + new[k] - old[k]

    # ApplyRandomRotation monte_carlo_utils.f90:30-66
    def ApplyRandomRotation(self, rng, off, full):
        na = off.shape[0]
        if na == 1:
            return off.copy()
        theta = rng() * TWOPI if full else (rng() - 0.5) * self.rotation_step
        axis = int(rng() * 3.0) + 1
        c, s = math.cos(theta), math.sin(theta)
        R = np.eye(3)
        if axis == 1:
            R[1, 1], R[1, 2], R[2, 1], R[2, 2] = c, -s, s, c
        elif axis == 2:
            R[0, 0], R[0, 2], R[2, 0], R[2, 2] = c, s, -s, c
        else:
            R[0, 0], R[0, 1], R[1, 0], R[1, 1] = c, -s, s, c
        out = np.empty_like(off)
        for a in range(na):
            v = off[a]
            for i in range(3):
                out[a, i] = R[i, 0] * v[0] + R[i, 1] * v[1] + R[i, 2] * v[2]
        return out

    def _trial(self, rng, kind, r, m, com, off, prob, r2=-1):
        """energies, acceptance probability and decision of one move; `prob(old, new)` is the host's acceptance rule"""
        if self.fused:
            u = rng()
            return self.gpu.trial_step(kind, r, m, com, off, u=u, res_new=r2, chain=self.chain)
        if kind == MMC_SWAP:
            old, new = self.gpu.trial_energy_swap(r, m, r2, com, off, chain=self.chain)
        else:
            old, new = self.gpu.trial_energy(kind, r, m, com, off, chain=self.chain)
        p = prob(old, new)
        return old, new, p, rng() <= p

    def _commit(self, acc):
        if not self.fused:
            self.gpu.trial_commit(acc, chain=self.chain)

    def _record(self, kind, r, m, acc, old, new, p):
        rec = np.zeros(1, dtype=RECORD_DTYPE)[0]
        rec["kind"], rec["res"], rec["mol"], rec["accepted"], rec["prob"] = kind, r, m, acc, p
        rec["old_e"], rec["new_e"] = old, new
        return rec

    # tilted cell a = (lx,0,0), b = (xy,ly,0), c = (xz,yz,lz): same arithmetic as the library and the CPU restatement
    @property
    def tri(self):
        return bool(np.any(np.asarray(self.sys.tilt) != 0.0))

    def frac(self, p):
        L, (xy, xz, yz) = self.sys.box_len, self.sys.tilt
        sz = (1.0 / L[2]) * p[2]
        sy = (1.0 / L[1]) * (p[1] - yz * sz)
        sx = (1.0 / L[0]) * (p[0] - xy * sy - xz * sz)
        return [sx, sy, sz]

    def cart(self, s):
        L, (xy, xz, yz) = self.sys.box_len, self.sys.tilt
        return np.array([L[0] * s[0] + xy * s[1] + xz * s[2], L[1] * s[1] + yz * s[2], L[2] * s[2]])

    def Translation(self, rng, r, m):
        if m < 0:
            return None
        self.counter["trial_translations"] += 1
        u = [rng() - 0.5 for _ in range(3)]
        com = self.com[r][m].copy()
        if self.tri:      # ApplyPBC, triclinic branch: s = (a b c)^-1 (r - r0), s mod 1, r = r0 + (a b c) s
            s = self.frac(np.array([com[d] + u[d] * self.translation_step for d in range(3)]) - self.sys.box_lo)
            s = [v - math.floor(v) for v in s]
            com = self.sys.box_lo + self.cart([0.0 if v >= 1.0 else v for v in s])
        else:
            for d in range(3):
                x = com[d] + u[d] * self.translation_step
                com[d] = self.sys.box_lo[d] + f_modulo(x - self.sys.box_lo[d], self.sys.box_len[d])  # ApplyPBC
        old, new, p, acc = self._trial(rng, MMC_TRANSLATION, r, m, com, self.off[r][m],
                                       lambda o, n: self.mc_acceptance_probability(o, n, r, MMC_TRANSLATION))
        if acc:
            self._accept_move(old, new); self.counter["translations"] += 1; self.com[r][m] = com
        self._commit(acc)
        return self._record(MMC_TRANSLATION, r, m, acc, old, new, p)

    def Rotation(self, rng, r, m):
        if self.sys.natoms[r] == 1 or m < 0:
            return None
        self.counter["trial_rotations"] += 1
        off = self.ApplyRandomRotation(rng, self.off[r][m], full=False)
        old, new, p, acc = self._trial(rng, MMC_ROTATION, r, m, self.com[r][m], off,
                                       lambda o, n: self.mc_acceptance_probability(o, n, r, MMC_ROTATION))
        if acc:
            self._accept_move(old, new); self.counter["rotations"] += 1; self.off[r][m] = off
        self._commit(acc)
        return self._record(MMC_ROTATION, r, m, acc, old, new, p)

    def CreateMolecule(self, rng, r, m):
        if m >= self.sys.capacity[r]:
            raise RuntimeError("molecule index exceeds NB_MAX_MOLECULE")  # CheckMoleculeIndex
        self.counter["trial_creations"] += 1
        self.count[r] += 1
        if self.tri:      # bounds(:,1) + matmul(matrix, u), cell vectors as columns
            com = self.sys.box_lo + self.cart([rng() for _ in range(3)])
        else:
            com = np.array([self.sys.box_lo[d] + self.sys.box_len[d] * rng() for d in range(3)])
        rsv = -1
        if self.has_reservoir:   # InsertAndOrientMolecule create_molecule.f90:185-193
            rsv = int(rng() * len(self.rsv[r])); off = self.rsv[r][rsv].copy()
        else:
            off = self.ApplyRandomRotation(rng, self.off[r][0], full=True)  # geometry of slot 1 (Q9)
        self.com[r][m] = com; self.off[r][m] = off  # written at proposal time, kept on reject
        old, new, p, acc = self._trial(rng, MMC_CREATION, r, m, com, off,
                                       lambda o, n: self.mc_acceptance_probability(o, n, r, MMC_CREATION))
        if acc:
            self._accept_insdel(old, new); self.counter["creations"] += 1
            if self.has_reservoir and self.rsv[r]:   # AcceptCreationMove :118-130: pool slot <- last pool slot
                self.rsv[r][rsv] = self.rsv[r][-1]; self.rsv[r].pop()
        else:
            self.count[r] -= 1
        self._commit(acc)
        return self._record(MMC_CREATION, r, m, acc, old, new, p)

    def DeleteMolecule(self, rng, r, m):
        if self.count[r] == 0:
            return None
        self.counter["trial_deletions"] += 1
        last = self.count[r] - 1
        self.count[r] -= 1          # N is read after the count update (monte_carlo_utils.f90:199-214)
        old, new, p, acc = self._trial(rng, MMC_DELETION, r, m, None, None,
                                       lambda o, n: self.mc_acceptance_probability(o, n, r, MMC_DELETION))
        if acc:
            self._accept_insdel(old, new); self.counter["deletions"] += 1
            if self.has_reservoir:   # AcceptDeletionMove delete_molecule.f90:148-166 (appends slot `last`, three placement draws)
                rng(); rng(); rng()
                self.rsv[r].append(self.off[r][last].copy())
            self.com[r][m] = self.com[r][last]; self.off[r][m] = self.off[r][last]  # RemoveMolecule
        else:
            self.count[r] += 1
        self._commit(acc)
        return self._record(MMC_DELETION, r, m, acc, old, new, p)

    # mc_acceptance_probability_swap monte_carlo_utils.f90:228-268 (counts before the move)
    def mc_acceptance_probability_swap(self, old, new, type_old, type_new):
        de = new[E_TOTAL] - old[E_TOTAL]
        try:
            ex = math.exp(-de / self.sys.temperature)
        except OverflowError:
            ex = math.inf
        comb = float(self.count[type_old]) / (float(self.count[type_new]) + 1.0)
        return min(1.0, (float(self.sys.fugacity[type_new]) / float(self.sys.fugacity[type_old])) * comb * ex)

    def SwapMolecule(self, rng, r, m):
        """Identity exchange: the reference ships only the acceptance rule above and never calls the move
        (monte_carlo.f90:50-75); the move is composed from its deletion and creation steps: remove (r, m), insert a
        molecule of another active type at the same centre of mass (slot-1 template, random orientation)."""
        s = self.sys
        active = [i for i in range(s.nres) if s.is_active[i] == 1]
        if len(active) < 2 or m < 0:
            return None
        pick = min(int(rng() * (len(active) - 1)), len(active) - 2) if len(active) > 2 else 0
        r2 = [i for i in active if i != r][pick]
        m2 = int(self.count[r2])
        if m2 >= s.capacity[r2]:
            raise RuntimeError("molecule index exceeds NB_MAX_MOLECULE")
        self.counter["trial_swaps"] += 1
        com = self.com[r][m].copy()
        off = self.ApplyRandomRotation(rng, self.off[r2][0], full=True)
        self.com[r2][m2] = com; self.off[r2][m2] = off
        old, new, p, acc = self._trial(rng, MMC_SWAP, r, m, com, off, lambda o, n: self.mc_acceptance_probability_swap(o, n, r, r2), r2=r2)
        if acc:
            self._accept_insdel(old, new); self.counter["swaps"] += 1
            last = self.count[r] - 1
            self.com[r][m] = self.com[r][last]; self.off[r][m] = self.off[r][last]
            self.count[r] -= 1; self.count[r2] += 1
        self._commit(acc)
        return self._record(MMC_SWAP, r + 256 * r2, m, acc, old, new, p)

    def AdjustMoveStepSizes(self):  # monte_carlo_utils.f90:99-129 (single-precision ratios, Q6)
        if not self.recalibrate:
            return
        c = self.counter
        if c["trial_translations"] > MIN_TRIALS_FOR_RECALIBRATION:
            acc = float(np.float32(c["translations"]) / np.float32(c["trial_translations"]))
            if acc - TARGET_ACCEPTANCE > TOL_ACCEPTANCE:
                self.translation_step = min(self.translation_step * 1.05, MAX_TRANSLATION_STEP)
            elif acc - TARGET_ACCEPTANCE < TOL_ACCEPTANCE:
                self.translation_step = max(self.translation_step * 0.95, MIN_TRANSLATION_STEP)
        if c["trial_rotations"] > MIN_TRIALS_FOR_RECALIBRATION:
            acc = float(np.float32(c["rotations"]) / np.float32(c["trial_rotations"]))
            if acc - TARGET_ACCEPTANCE > TOL_ACCEPTANCE:
                self.rotation_step = min(self.rotation_step * 1.05, MAX_ROTATION_ANGLE)
            elif acc - TARGET_ACCEPTANCE < TOL_ACCEPTANCE:
                self.rotation_step = min(self.rotation_step * 1.95, MIN_ROTATION_ANGLE)

    def step(self, rng):
        s = self.sys
        active = [i for i in range(s.nres) if s.is_active[i] == 1]
        r = active[int(rng() * len(active))]                       # PickRandomResidueType
        m = -1
        if self.count[r] != 0:                                     # PickRandomMoleculeIndex
            m = min(int(rng() * self.count[r]), self.count[r] - 1)
        draw = rng()
        if draw <= s.p_translation:
            return self.Translation(rng, r, m)
        if draw <= s.p_rotation + s.p_translation:
            return self.Rotation(rng, r, m)
        if draw <= s.p_rotation + s.p_translation + s.p_swap:
            return self.SwapMolecule(rng, r, m)
        if rng() <= PROB_CREATE_DELETE:
            return self.CreateMolecule(rng, r, int(self.count[r]))
        return self.DeleteMolecule(rng, r, m)

    def MonteCarloLoop(self, rng, nb_block, nb_step):
        recs = []
        for _ in range(nb_block):
            for _ in range(nb_step):
                rec = self.step(rng)
                recs.append(rec if rec is not None else np.zeros(1, dtype=RECORD_DTYPE)[0])
            self.AdjustMoveStepSizes()
        return np.array(recs, dtype=RECORD_DTYPE)
