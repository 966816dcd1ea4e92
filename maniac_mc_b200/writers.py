"""Per-block output files in the reference's formats, one directory per chain.

The reference writes, after every block (`monte_carlo.f90:82-84` -> `UpdateFiles`, `write_utils.f90:418-434`):
`energy.dat`, `number_<RESNAME>.dat`, `moves.dat` (`WriteEnergyAndCount`, `write_utils.f90:96-188`) and
`trajectory.lammpstrj` (`WriteLAMMPSTRJ`, `write_utils.f90:12-94`), all under ONE global `output_path`
(`simulation_state.f90:9`).  An isotherm sweep runs many chains per process, so every chain gets its own
directory `<root>/p<point>_r<replica>/` holding the same files, byte-compatible with the reference's Fortran edit
descriptors (I10, F16.6, I12, F15.8, F12.7 ...), quirks included:

* `moves.dat` has 9 columns for 11 header names, and its last column repeats `counter%deletions` where the header
  says `Rotate_Moves` (`write_utils.f90:182-186`); `literal=False` writes the accepted rotations instead;
* `number_<RES>.dat` gets a row only on blocks where the count is non-zero (`write_utils.f90:150`);
* the trajectory's TIMESTEP line is `input%nb_block`, the TOTAL number of blocks, on every frame
  (`write_utils.f90:44`); `literal=False` writes the current block;
* the trajectory's box bounds are -L/2 .. L/2 whatever the data file's `xlo` was (`write_utils.f90:48-50`).

Host-side only (numpy + files): the device side of the block exchange is `mmc_snapshot_begin/_wait`, which lets the
files of block b be written while the GPU runs block b+1 (`BlockPipeline`).
"""
from __future__ import annotations

import os

import numpy as np

KB_KCALMOL = 0.0019872041          # constants.f90:11

ENERGY_HEADER = ("#    block        total        recipCoulomb"
                 "     non-coulomb      coulomb     ewald_self    intramolecular-coulomb")
NUMBER_HEADER = "# Block   Active_Molecules"
MOVES_HEADER = ("# Block   Trial_Trans   Trans_Moves   Trial_Create   Create_Moves   "
                "Trial_Delete   Delete_Moves   Trial_Rotate   Rotate_Moves   "
                "Trial_BigMove   Big_Moves")


def _fint(v, w):
    """Fortran Iw: right-justified, a field of asterisks on overflow"""
    s = str(int(v))
    return s.rjust(w) if len(s) <= w else "*" * w


def _ffix(v, w, d):
    """Fortran Fw.d as gfortran prints it (leading zero kept when it fits, asterisks on overflow)"""
    s = f"{float(v):.{d}f}"
    if len(s) > w and s.startswith("0."):
        s = s[1:]
    elif len(s) > w and s.startswith("-0."):
        s = "-" + s[2:]
    return s.rjust(w) if len(s) <= w else "*" * w


def _nint(x):
    """Fortran nint: halves round away from zero (numpy's rint rounds them to even)"""
    x = np.asarray(x, dtype=np.float64)
    return np.trunc(x + np.copysign(0.5, x))


def wrap_into_box(pos, box_len):
    """WrapIntoBox for box types 1 and 2 (`geometry_utils.f90:247-253`): x - L * nint(x / L)"""
    pos = np.asarray(pos, dtype=np.float64)
    return pos - box_len * _nint(pos / box_len)


def energy_row(block, e):
    """e = (non_coulomb, coulomb, recip_coulomb, ewald_self, intra_coulomb, total) in Kelvin, the C ABI's order"""
    k = [x * KB_KCALMOL for x in e]
    cols = [k[5], k[2], k[0], k[1], k[3], k[4]]           # total, recip, non-coulomb, coulomb, self, intra
    return _fint(block, 10) + "".join(" " + _ffix(c, 16, 6) for c in cols)


def number_row(block, n):
    return _fint(block, 10) + " " + _fint(n, 10)


# field order of `mmc_counters` (include/maniac_gpu.h), the Counters ctypes class and the snapshot rows
COUNTER_FIELDS = ("translations", "rotations", "creations", "deletions", "trial_translations", "trial_rotations",
                  "trial_creations", "trial_deletions", "swaps", "trial_swaps")


def moves_row(block, ctr, literal=True):
    """One `moves.dat` row (`write_utils.f90:178-186`).  `ctr` is one chain's counters in `mmc_counters` STRUCT order
    (COUNTER_FIELDS: the four accepted counts, then the four trial counts, then swaps, trial_swaps) or a mapping by
    field name.  Columns: block, trial_translations, translations, trial_creations, creations, trial_deletions,
    deletions, trial_rotations and -- literally as the reference writes it (:186) -- deletions again in the
    Rotate_Moves column; `literal=False` puts the accepted rotations there."""
    c = ctr if isinstance(ctr, dict) else dict(zip(COUNTER_FIELDS, ctr))
    last = c["deletions"] if literal else c["rotations"]
    cols = [block, c["trial_translations"], c["translations"], c["trial_creations"], c["creations"],
            c["trial_deletions"], c["deletions"], c["trial_rotations"], last]
    return " ".join(_fint(int(v), 12) for v in cols)


def lammpstrj_frame(timestep, box_len, types, is_active, natoms, count, com, offsets):
    """One frame of WriteLAMMPSTRJ.  `com` [sum(count), 3] and `offsets` [sum(count*natoms), 3] are packed in
    (residue, molecule[, atom]) order, `types` is [nres, max_atoms] (1-based atom types)."""
    n_atoms = int((np.asarray(count) * np.asarray(natoms)).sum())
    out = ["ITEM: TIMESTEP", _fint(timestep, 10), "ITEM: NUMBER OF ATOMS", _fint(n_atoms, 10),
           "ITEM: BOX BOUNDS pp pp pp"]
    for d in range(3):
        out.append(_ffix(-box_len[d] / 2, 15, 8) + " " + _ffix(box_len[d] / 2, 15, 8))
    out.append("ITEM: ATOMS id type x y z")
    atom_id = 0
    im = ia = 0
    for r in range(len(natoms)):
        nm, na = int(count[r]), int(natoms[r])
        if nm == 0:
            continue
        c = np.asarray(com[im:im + nm], dtype=np.float64)
        o = np.asarray(offsets[ia:ia + nm * na], dtype=np.float64).reshape(nm, na, 3)
        if is_active[r]:
            pos = wrap_into_box(c, box_len)[:, None, :] + o      # molecules stay whole: only the centre is wrapped
        else:
            pos = wrap_into_box(c[:, None, :] + o, box_len)      # framework: every atom wrapped
        for m in range(nm):
            for a in range(na):
                atom_id += 1
                p = pos[m, a]
                out.append(_fint(atom_id, 6) + " " + _fint(types[r][a], 4) + "".join(" " + _ffix(x, 12, 7) for x in p))
        im += nm
        ia += nm * na
    return "\n".join(out) + "\n"


class ChainWriter:
    """The reference's per-block files for ONE chain in its own directory."""

    def __init__(self, directory, system, literal=True, nb_block=0, trajectory=True):
        self.dir, self.sys, self.literal, self.nb_block, self.trajectory = directory, system, literal, nb_block, trajectory
        os.makedirs(directory, exist_ok=True)
        self.names = list(system.names) if system.names else [f"res{r + 1}" for r in range(system.nres)]

    def _append(self, name, header, row, block):
        path = os.path.join(self.dir, name)
        fresh = block == 0 or not os.path.exists(path)           # block 0: STATUS='REPLACE'
        with open(path, "w" if block == 0 else "a") as f:
            if fresh:
                f.write(header + "\n")
            f.write(row + "\n")

    def write_block(self, block, count, com, offsets, energies, counters):
        """`count` [nres]; `com` / `offsets` packed over ALL residues (framework included) as in `System`;
        `energies` [6] Kelvin; `counters` in `mmc_counters` order."""
        s = self.sys
        self._append("energy.dat", ENERGY_HEADER, energy_row(block, energies), block)
        for r in range(s.nres):
            if count[r] != 0 and s.is_active[r] == 1:
                self._append(f"number_{self.names[r]}.dat", NUMBER_HEADER, number_row(block, count[r]), block)
        self._append("moves.dat", MOVES_HEADER, moves_row(block, counters, self.literal), block)
        if self.trajectory:
            ts = self.nb_block if self.literal else block
            frame = lammpstrj_frame(ts, s.box_len, s.types, s.is_active, s.natoms, count, com, offsets)
            with open(os.path.join(self.dir, "trajectory.lammpstrj"), "a" if block > 0 else "w") as f:
                f.write(frame)


def chain_directory(root, point, replica):
    """per-chain output directory of an isotherm sweep: <root>/p<point>_r<replica>/"""
    return os.path.join(root, f"p{point:03d}_r{replica:02d}")


def unpack_chain(system, count, gcom, goff):
    """One chain of a snapshot (`count` [nres], `gcom` [g_mols, 3], `goff` [g_atoms, 3]: fixed-capacity slots of the
    ACTIVE residues, see include/maniac_gpu.h) -> packed (com, offsets) over all residues; the inactive residues'
    coordinates never leave the host and come from `system`."""
    mo0, ao0 = system.mol_offsets()
    coms, offs = [], []
    mb = ab = 0
    for r in range(system.nres):
        na = int(system.natoms[r])
        if system.is_active[r] == 1:
            n = int(count[r])
            coms.append(gcom[mb:mb + n])
            offs.append(goff[ab:ab + n * na])
            mb += int(system.capacity[r])
            ab += int(system.capacity[r]) * na
        else:
            n = int(system.count[r])
            coms.append(system.com[mo0[r]:mo0[r] + n])
            offs.append(system.offsets[ao0[r]:ao0[r] + n * na])
    return np.concatenate(coms), np.concatenate(offs)


class BlockPipeline:
    """Blocks of MC steps on every chain of one handle with the reference's per-block files written WHILE the next
    block runs: block b is queued, its snapshot is queued behind it (`mmc_snapshot_begin`, slot b mod 2), block b+1
    is queued at once, and only then does the host wait for snapshot b and write its files.

    `gpu` is a `ManiacGPU`; `writers[c]` a `ChainWriter` (or None to skip the chain); `on_block(block, snapshot)` is
    called with every snapshot (count, com, offsets, energies, counters) after its files are written."""

    def __init__(self, gpu, writers, on_block=None, recalibrate=False):
        # recalibrate: AdjustMoveStepSizes after every block (monte_carlo.f90:78, input%recalibrate_moves); the call is
        # stream-ordered on the device, so it does not stall the pipeline
        self.gpu, self.writers, self.on_block, self.recalibrate = gpu, writers, on_block, recalibrate
        self.block = 0
        self.last = None

    def _drain(self, slot, block):
        cnt, gcom, goff, e, ctr = self.gpu.snapshot_wait(slot)
        self.last = (cnt, gcom, goff, e, ctr)
        for c, w in enumerate(self.writers):
            if w is None:
                continue
            com, off = unpack_chain(w.sys, cnt[c], gcom[c], goff[c])
            w.write_block(block, cnt[c], com, off, e[c], ctr[c])
        if self.on_block is not None:
            self.on_block(block, self.last)

    def write_initial(self):
        """block 0 row of every file (the reference writes it before the first step, `monte_carlo.f90:37`)"""
        self.gpu.snapshot_begin(0)
        self._drain(0, 0)

    def run(self, n_blocks, n_steps, seed=0, pipelined=True):
        """`pipelined=False` waits for every snapshot and writes its files before the next block is queued (what the
        reference's loop does); only there to measure what the overlap buys"""
        pending = None
        for _ in range(n_blocks):
            self.block += 1
            self.gpu.run_batch_async(n_steps, seed=seed)
            if self.recalibrate:
                self.gpu.adjust_step_sizes()
            slot = self.block & 1
            self.gpu.snapshot_begin(slot)
            if not pipelined:
                self._drain(slot, self.block)
                continue
            if pending is not None:
                self._drain(*pending)                            # files of the previous block, GPU busy with this one
            pending = (slot, self.block)
        if pending is not None:
            self._drain(*pending)
        return self.last
