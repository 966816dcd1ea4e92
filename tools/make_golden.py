"""Converts the reference's shipped fixture systems (mc-topology submodule) into the packed
``System`` .npz files under tests/golden/, and records the reference's own known answers.

Run in the build container only (needs /root/reference); the GPU box uses the committed files.
    python tools/make_golden.py
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from maniac_mc_b200 import builders  # noqa: E402

REF = "/root/reference/mc-topology"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

CASES = {
    # name: (dir, maniac, data, inc, capacity)
    "lj_gas": ("testcase-energy/LJ-gas", "input.maniac", "topology.data", "parameters.inc", 256),
    "h2o_gas": ("testcase-energy/H2O-gas", "input.maniac", "topology.data", "parameters.inc", 512),
    "h2o_ads": ("testcase-adsorption/H2O", "input.maniac", "topology.data", "parameters.inc", 512),
    "zif8_h2o": ("testcase-adsorption/ZIF8-H2O", "zif8-water.maniac", "zif8-water.data", "zif8-water.inc", 512),
    # two active species (methanol 6 sites, water 4 sites) in ZIF-8, swap_proba 0.2: the swap-move fixture
    # non-cubic orthorhombic zeolite cell (40.18 x 39.48 x 26.28 A -> kmax (11, 10, 7), K = 1697) + linear 3-site CO2
    "mfi_co2": ("testcase-adsorption/MFI-CO2", "MFI-CO2.maniac", "MFI-CO2.data", "MFI-CO2.inc", 256),
    # slit pore: 1912-atom wall + single-site fluid (no rotation moves, box origin at -20)
    "slit": ("testcase-adsorption/SLIT", "input.maniac", "topology.data", "parameters.inc", 256),
    "zif8_ch4o_h2o": ("testcase-adsorption/ZIF8-CH4O-H2O", "input.maniac", "topology.data", "parameters.inc", 256),
    # the reference's only tilted cell: 71 two-site dipoles, 40 A box with xy xz yz = -20 -20 20 (despite the directory name)
    "dipole_tri": ("testcase-energy/DIPOLE-orthorhombic", "dipole.maniac", "dipole.data", "dipole.inc", 256),
}

# Known answers held by the reference's own tests (kcal/mol):
#   tests/integration/energy/LJ-gas/run-test.sh:18 parses TotEng from
#   mc-topology/testcase-energy/LJ-gas/log.lammps:79-80  ("0 -38.589253 -0.40707303 49.708669 -87.890849")
#   tests/integration/energy/H2O-gas/run-test.sh:15-16   (ref_energy=-32.822927, tolerance=0.01)
#   mc-topology/testcase-adsorption/DIPOLE-orthorhombic/log.lammps:115: step 0 of LAMMPS' own GCMC run of the TILTED cell
#   (triclinic box (-20 -20 -20) to (20 20 20) with tilt (-20 -20 20), 142 atoms = 71 dipoles, ewald 1e-5, kmax 6 6 6,
#   gewald 0.1622) "0  -2.8812361  -1.2645192  40.097413  -41.71413  71"; its starting configuration is the 142-atom
#   testcase-energy/DIPOLE-orthorhombic/dipole.data (fixture dipole_tri).  No reference script asserts it; the tolerance
#   is the one the reference's energy tests use (run-test.sh: 0.01 kcal/mol)
#   mc-topology/testcase-adsorption/H2O/number.dat:3-103: N(H2O) of LAMMPS fix gcmc at 0.5 atm, 300 K, 101 samples
KAT = {
    "lj_gas": {"TotEng": -38.589253, "E_vdwl": -0.40707303, "E_coul": 49.708669, "E_long": -87.890849, "tolerance": 0.01},
    "h2o_gas": {"TotEng": -32.822927, "tolerance": 0.01},
    "dipole_tri": {"TotEng": -2.8812361, "E_vdwl": -1.2645192, "E_coul": 40.097413, "E_long": -41.71413, "N": 71, "tolerance": 0.01},
}


def lammps_number_series(path):
    """second column of a LAMMPS `fix ave/time`-style number.dat (comment lines skipped)"""
    out = []
    for ln in open(path):
        f = ln.split()
        if not f or f[0].startswith("#"):
            continue
        out.append(float(f[1]))
    return out


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    for name, (d, m, t, p, cap) in CASES.items():
        s = builders.from_files(*(os.path.join(REF, d, f) for f in (m, t, p)), capacity=cap)
        s.save(os.path.join(OUT, name + ".npz"))
        print(name, "counts", s.count.tolist(), "natoms", s.natoms.tolist(), "alpha %.6f" % s.alpha, "kmax", s.kmax,
              "K", s.num_kvectors, "L", s.box_len.tolist())
    KAT["h2o_ads_gcmc"] = {"source": "mc-topology/testcase-adsorption/H2O/number.dat:3-103 (LAMMPS fix gcmc, 0.5 atm, 300 K)",
                           "N": lammps_number_series(os.path.join(REF, "testcase-adsorption/H2O/number.dat"))}
    with open(os.path.join(OUT, "reference_kat.json"), "w") as fh:
        json.dump(KAT, fh, indent=1)
