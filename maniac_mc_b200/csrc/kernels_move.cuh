// kernels_move.cuh -- the single-thread parts of one trial move: selection + proposal geometry (propose_move),
// the Trial descriptor (build_trial), energy assembly and acceptance probability (finalize_trial), commit
// (commit_trial), plus the small strict-mode / k-sharded kernels around them.  The sweeps are in kernels_sweep.cuh.
#pragma once
#include "common.cuh"

// ----------------------------------------------------------------------------------------------
// small helpers on chain-local guest storage
__device__ __forceinline__ double* g_com(const DevConst& dc, const ChainState& cs, int c, int r, int m) {
    return cs.gcom + ((size_t)c * dc.g_mols + dc.g_mol_base[r] + m) * 3;
}
__device__ __forceinline__ size_t g_atom(const DevConst& dc, int c, int r, int m, int a) {
    return (size_t)c * dc.g_atoms + dc.g_atom_base[r] + (size_t)m * dc.natoms[r] + a;
}

#define MMC_PRE_DRAWS 16   // uniforms of one MC step computed ahead by 16 lanes (a step draws at most 9 + 1 + 3)
struct Rng {
    const RngSrc* src; ChainState cs; int c; int draw;
    const double* pre = nullptr;   // Philox draws 0 .. MMC_PRE_DRAWS-1 of this step, computed in parallel (same values as below)
    __device__ double next() {
        if (src->tape) {
            // a tape shorter than the draws consumed must not run into the next chain's tape (a host-side replay aborts on
            // exhaustion): flag it (MMC_ERR_TAPE, reported by the host entry point) and return a harmless uniform
            const long long p = cs.tape_pos[c]++;
            if (p >= src->tape_stride) { atomicOr(cs.error_flag, MMC_ERRF_TAPE); return 0.5; }
            return src->tape[(size_t)c * src->tape_stride + p];
        }
        if (pre && draw < MMC_PRE_DRAWS) return pre[draw++];
        return philox_uniform(src->seed, cs.chain_id[c], src->step, (uint32_t)draw++);
    }
};

// RotationMatrix helper_utils.f90:39-77 + matmul of ApplyRandomRotation monte_carlo_utils.f90:63-64
__device__ inline void rotate_offsets(int axis, double theta, int nat, const double (*in)[3], double (*out)[3]) {
    double s, c; sincos(theta, &s, &c);
    double R[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    if (axis == 1) { R[1][1] = c; R[1][2] = -s; R[2][1] = s; R[2][2] = c; }
    else if (axis == 2) { R[0][0] = c; R[0][2] = s; R[2][0] = -s; R[2][2] = c; }
    else { R[0][0] = c; R[0][1] = -s; R[1][0] = s; R[1][1] = c; }
    for (int a = 0; a < nat; a++) {
        double v0 = in[a][0], v1 = in[a][1], v2 = in[a][2];
        for (int i = 0; i < 3; i++) out[a][i] = R[i][0] * v0 + R[i][1] * v1 + R[i][2] * v2;
    }
}

// Fills the Trial of chain c for a move of `kind` on (res, mol) with proposed geometry
// com_new/off_new (ignored for deletion).  Mirrors which sweeps ComputeOldEnergy /
// ComputeNewEnergy (monte_carlo_utils.f90:347,275) request for each move type.  Single thread.
// `t` is the caller's Trial: shared memory in the batched kernel and in the fused strict-mode kernel, cs.trial[c] (global) on
// the k-sharded path.  write_slot: also write the proposed geometry of a creation / swap into its slot (done by ONE CTA when
// several build the same trial).
__device__ inline void build_trial(const DevConst& dc, const ChainState& cs, int c, Trial& t, int kind, int r, int m,
                                   const double* com_new, const double (*off_new)[3], int r2 = -1, bool write_slot = true,
                                   const double2* fw_pre = nullptr) {
    const int nat = dc.natoms[r];
    if (kind != MMC_SWAP) r2 = r;
    const int nat2 = dc.natoms[r2];
    t.kind = kind; t.res = r; t.mol = m; t.nat = nat; t.res2 = r2; t.nat2 = nat2; t.accepted = 0; t.prob = 0.0; t.rsv = -1;
    t.fw_new_ok = 0; t.fw_old_fresh = 0;
    // issued first, consumed last: the stores to `t` below do not have to wait for it
    // (fw_pre: the caller loaded the entry of (r, m) already, together with the molecule's geometry)
    double2 fwv = make_double2(__longlong_as_double(-1LL), 0.0);
    if (dc.fw_cache && (kind == MMC_TRANSLATION || kind == MMC_ROTATION || kind == MMC_DELETION))
        fwv = fw_pre ? *fw_pre : cs.fw_pair[(size_t)c * dc.g_mols + dc.g_mol_base[r] + m];
    for (int a = 0; a < nat; a++) { t.q[a] = dc.q[r][a]; t.type[a] = dc.type[r][a]; }
    for (int a = 0; a < nat2; a++) { t.q2[a] = dc.q[r2][a]; t.type2[a] = dc.type[r2][a]; }
    const int cnt = cs.count[c * MMC_MAX_RES + r];
    if (kind != MMC_CREATION) {  // old geometry of the selected slot
        const double* com = g_com(dc, cs, c, r, m);
        for (int a = 0; a < nat; a++) {
            const double* off = cs.goff + g_atom(dc, c, r, m, a) * 3;
            for (int d = 0; d < 3; d++) t.p_old[a][d] = com[d] + off[d];
        }
    }
    if (kind != MMC_DELETION) {
        for (int d = 0; d < 3; d++) t.com_new[d] = com_new[d];
        for (int a = 0; a < nat2; a++)
            for (int d = 0; d < 3; d++) {
                t.off_new[a][d] = off_new[a][d];
                t.p_new[a][d] = com_new[d] + off_new[a][d];
                t.p_plus[a][d] = t.p_new[a][d];
            }
    }
    switch (kind) {
        case MMC_TRANSLATION: case MMC_ROTATION:
            t.do_old = 1; t.do_new = 1; t.r_plus = 1; t.r_minus = 1; t.excl = 1; break;
        case MMC_CREATION:
            // the new molecule sits in slot `cnt`, outside the swept range, so nothing to exclude
            t.do_old = 0; t.do_new = 1; t.r_plus = 1; t.r_minus = 0; t.excl = 0;
            // like the reference (create_molecule.f90:180-199) the geometry is written into the slot at
            // proposal time and stays there even if the move is rejected
            if (write_slot) {
                double* com = g_com(dc, cs, c, r, m);
                for (int d = 0; d < 3; d++) com[d] = com_new[d];
                for (int a = 0; a < nat; a++) {
                    size_t ia = g_atom(dc, c, r, m, a);
                    for (int d = 0; d < 3; d++) cs.goff[ia * 3 + d] = off_new[a][d];
                    cs.gx[ia] = t.p_new[a][0]; cs.gy[ia] = t.p_new[a][1]; cs.gz[ia] = t.p_new[a][2];
                }
            }
            break;
        case MMC_SWAP:
            // removed molecule (r, m): old sweep with its own atoms, excluded from both sweeps; inserted molecule of type
            // r2 in slot count(r2), outside the swept range; like a creation its geometry is written at proposal time
            t.do_old = 1; t.do_new = 1; t.r_plus = 1; t.r_minus = 1; t.excl = 1;
            if (write_slot) {
                const int m2 = cs.count[c * MMC_MAX_RES + r2];
                double* com = g_com(dc, cs, c, r2, m2);
                for (int d = 0; d < 3; d++) com[d] = com_new[d];
                for (int a = 0; a < nat2; a++) {
                    size_t ia = g_atom(dc, c, r2, m2, a);
                    for (int d = 0; d < 3; d++) cs.goff[ia * 3 + d] = off_new[a][d];
                    cs.gx[ia] = t.p_new[a][0]; cs.gy[ia] = t.p_new[a][1]; cs.gz[ia] = t.p_new[a][2];
                }
            }
            break;
        case MMC_DELETION:
            t.do_old = 1; t.do_new = 0; t.excl = 1;
            if (dc.deletion_mode == MMC_DELETION_REFERENCE) {
                // Q2: ComputeNewEnergy passes is_creation=deletion_flag (monte_carlo_utils.f90:308), so the
                // phases of the molecule now in this slot (the former last one, delete_molecule.f90:107-114)
                // are ADDED
                t.r_plus = 1; t.r_minus = 0;
                const int last = cnt - 1;
                const double* com = g_com(dc, cs, c, r, last);
                for (int a = 0; a < nat; a++) {
                    const double* off = cs.goff + g_atom(dc, c, r, last, a) * 3;
                    for (int d = 0; d < 3; d++) t.p_plus[a][d] = com[d] + off[d];
                }
            } else {  // physical: subtract the deleted molecule (is_deletion branch, ewald_energy.f90:246-250)
                t.r_plus = 0; t.r_minus = 1;
            }
            break;
    }
    t.fw_old[0] = fwv.x; t.fw_old[1] = fwv.y;
    t.fw_old_ok = !(isnan(fwv.x) || isnan(fwv.y));
}

// ----------------------------------------------------------------------------------------------
// strict mode: the host made the proposal (Fortran RandomTranslation / ApplyRandomRotation /
// InsertAndOrientMolecule); arguments by value, no H2D copy.
struct HostTrial { int chain, kind, res, mol, res2; double com[3]; double off[MMC_MAX_MOL_ATOMS][3]; };

__global__ void k_set_trial(DevConst dc, ChainState cs, HostTrial ht) {
    if (threadIdx.x == 0) build_trial(dc, cs, ht.chain, cs.trial[ht.chain], ht.kind, ht.res, ht.mol, ht.com, ht.off, ht.res2);
}

// batched mode: one step of MonteCarloLoop's selection logic (monte_carlo.f90:44-75) and the
// proposal geometry, draw order as in SURVEY 3.8.  Single thread per chain.
// lead: this caller performs the proposal's side effects (trial counters, the geometry of a creation / swap written into its
// slot).  The CTAs of a cluster that share one chain (k_chain_run_cl) all make the same proposal; only rank 0 leads.
__device__ inline void propose_move(const DevConst& dc, const ChainState& cs, const RngSrc& rs, int c, Trial& t, const double* pre = nullptr, bool lead = true) {
    t.kind = 0; t.error = 0; t.accepted = 0;
    for (int i = 0; i < MMC_NE; i++) { t.old_e[i] = 0.0; t.new_e[i] = 0.0; }
    Rng rng{&rs, cs, c, 0, pre};
    if (dc.n_active == 0) { t.draw = 0; return; }
    // PickRandomResidueType monte_carlo_utils.f90:136-164
    const int r = dc.active_list[(int)(rng.next() * dc.n_active)];
    const int cnt = cs.count[c * MMC_MAX_RES + r];
    const int nat = dc.natoms[r];
    // PickRandomMoleculeIndex :170-182
    int m = -1;
    if (cnt != 0) { m = (int)(rng.next() * cnt); if (m > cnt - 1) m = cnt - 1; }
    const double draw = rng.next();
    mmc_counters& ctr = cs.counters[c];
    double com_new[3]; double off_new[MMC_MAX_MOL_ATOMS][3];
    // the selected molecule's framework-energy cache entry: requested here, next to the loads of its geometry, used by build_trial
    double2 fwv = make_double2(__longlong_as_double(-1LL), 0.0);
    if (dc.fw_cache && m >= 0) fwv = cs.fw_pair[(size_t)c * dc.g_mols + dc.g_mol_base[r] + m];
    if (draw <= dc.p_translation) {
        if (m >= 0) {  // Translation translation.f90:36-112
            if (lead) ctr.trial_translations++;
            const double* com = g_com(dc, cs, c, r, m);
            const double step = cs.tstep[c];
            double u[3]; for (int d = 0; d < 3; d++) u[d] = rng.next() - 0.5;
            if (dc.tri) {   // ApplyPBC, triclinic branch (geometry_utils.f90:198-217): s = (a b c)^-1 (r - r0), s mod 1, back
                double rel[3], sf[3];
                for (int d = 0; d < 3; d++) rel[d] = (com[d] + u[d] * step) - dc.lo[d];
                for (int d = 2; d >= 0; d--) { sf[d] = frac_coord(dc, rel, d); sf[d] -= floor(sf[d]); if (sf[d] >= 1.0) sf[d] = 0.0; }
                com_new[0] = dc.lo[0] + (dc.L[0] * sf[0] + dc.tilt[0] * sf[1] + dc.tilt[1] * sf[2]);
                com_new[1] = dc.lo[1] + (dc.L[1] * sf[1] + dc.tilt[2] * sf[2]);
                com_new[2] = dc.lo[2] + dc.L[2] * sf[2];
            } else
            for (int d = 0; d < 3; d++) {
                double x = com[d] + u[d] * step;
                com_new[d] = dc.lo[d] + f_modulo(x - dc.lo[d], dc.L[d], dc.invL[d]);  // ApplyPBC
            }
            for (int a = 0; a < nat; a++) for (int d = 0; d < 3; d++) off_new[a][d] = cs.goff[g_atom(dc, c, r, m, a) * 3 + d];
            build_trial(dc, cs, c, t, MMC_TRANSLATION, r, m, com_new, off_new, -1, lead, &fwv);
        }
    } else if (draw <= dc.p_rotation + dc.p_translation) {
        if (nat != 1 && m >= 0) {  // Rotation rotation.f90:34-75
            if (lead) ctr.trial_rotations++;
            const double theta = (rng.next() - 0.5) * cs.rstep[c];  // ChooseRotationAngle
            const int axis = (int)(rng.next() * 3.0) + 1;
            double off_old[MMC_MAX_MOL_ATOMS][3];
            for (int a = 0; a < nat; a++) for (int d = 0; d < 3; d++) off_old[a][d] = cs.goff[g_atom(dc, c, r, m, a) * 3 + d];
            rotate_offsets(axis, theta, nat, off_old, off_new);
            const double* com = g_com(dc, cs, c, r, m);
            for (int d = 0; d < 3; d++) com_new[d] = com[d];
            build_trial(dc, cs, c, t, MMC_ROTATION, r, m, com_new, off_new, -1, lead, &fwv);
        }
    } else if (draw <= dc.p_rotation + dc.p_translation + dc.p_swap) {
        // swap (identity exchange): remove (r, m), insert a molecule of another active type at the same
        // centre of mass with that type's slot-1 template geometry in a random orientation (draws: type pick when more
        // than two active types, angle, axis)
        if (dc.n_active >= 2 && m >= 0) {
            int pick = dc.n_active > 2 ? (int)(rng.next() * (dc.n_active - 1)) : 0;
            if (pick > dc.n_active - 2) pick = dc.n_active - 2;
            int r2 = -1;
            for (int i = 0, k = 0; i < dc.n_active; i++) { const int rr = dc.active_list[i]; if (rr == r) continue; if (k == pick) { r2 = rr; break; } k++; }
            const int m2 = cs.count[c * MMC_MAX_RES + r2], nat2 = dc.natoms[r2];
            if (m2 >= dc.cap[r2]) { t.error = 1; }
            else {
                if (lead) ctr.trial_swaps++;
                const double* com = g_com(dc, cs, c, r, m);
                for (int d = 0; d < 3; d++) com_new[d] = com[d];
                double tmpl[MMC_MAX_MOL_ATOMS][3];
                for (int a = 0; a < nat2; a++) for (int d = 0; d < 3; d++) tmpl[a][d] = cs.goff[g_atom(dc, c, r2, 0, a) * 3 + d];
                if (nat2 != 1) {
                    const double theta = rng.next() * MMC_TWOPI;
                    const int axis = (int)(rng.next() * 3.0) + 1;
                    rotate_offsets(axis, theta, nat2, tmpl, off_new);
                } else for (int d = 0; d < 3; d++) off_new[0][d] = tmpl[0][d];
                build_trial(dc, cs, c, t, MMC_SWAP, r, m, com_new, off_new, r2, lead);
            }
        }
    } else {
        if (rng.next() <= MMC_PROB_CREATE_DELETE) {  // CreateMolecule create_molecule.f90:41-207
            m = cnt;
            if (m >= dc.cap[r]) { t.error = 1; }  // CheckMoleculeIndex -> AbortRun in the reference
            else {
                if (lead) ctr.trial_creations++;
                if (dc.tri) {   // bounds(:,1) + matmul(matrix, u) with the cell vectors as columns (create_molecule.f90:180-182)
                    double uu[3]; for (int d = 0; d < 3; d++) uu[d] = rng.next();
                    com_new[0] = dc.lo[0] + (dc.L[0] * uu[0] + dc.tilt[0] * uu[1] + dc.tilt[1] * uu[2]);
                    com_new[1] = dc.lo[1] + (dc.L[1] * uu[1] + dc.tilt[2] * uu[2]);
                    com_new[2] = dc.lo[2] + dc.L[2] * uu[2];
                } else
                for (int d = 0; d < 3; d++) com_new[d] = dc.lo[d] + dc.L[d] * rng.next();
                int rsv = -1;
                if (dc.has_reservoir) {  // InsertAndOrientMolecule create_molecule.f90:185-193: a random reservoir molecule's geometry, no rotation
                    rsv = (int)(rng.next() * cs.rcount[c * MMC_MAX_RES + r]);
                    const double* ro = cs.roff + ((size_t)c * dc.r_atoms + dc.r_atom_base[r] + (size_t)rsv * nat) * 3;
                    for (int a = 0; a < nat; a++) for (int d = 0; d < 3; d++) off_new[a][d] = ro[a * 3 + d];
                } else {
                    double tmpl[MMC_MAX_MOL_ATOMS][3];  // geometry of slot 1 of this type (Q9)
                    for (int a = 0; a < nat; a++) for (int d = 0; d < 3; d++) tmpl[a][d] = cs.goff[g_atom(dc, c, r, 0, a) * 3 + d];
                    if (nat != 1) {
                        const double theta = rng.next() * MMC_TWOPI;
                        const int axis = (int)(rng.next() * 3.0) + 1;
                        rotate_offsets(axis, theta, nat, tmpl, off_new);
                    } else for (int d = 0; d < 3; d++) off_new[0][d] = tmpl[0][d];
                }
                build_trial(dc, cs, c, t, MMC_CREATION, r, m, com_new, off_new, -1, lead);
                t.rsv = rsv;
            }
        } else if (cnt != 0) {  // DeleteMolecule delete_molecule.f90:41-199
            if (lead) ctr.trial_deletions++;
            build_trial(dc, cs, c, t, MMC_DELETION, r, m, nullptr, nullptr, -1, lead, &fwv);
        }
    }
    t.draw = rng.draw;
}

// ----------------------------------------------------------------------------------------------
// ComputeIntraResidueRealCoulombEnergySingleMol ewald_energy.f90:371-411 on explicit positions
// one pair of ComputeIntraResidueRealCoulombEnergySingleMol (ewald_energy.f90:371-411), without the 1/(4 pi eps0) factor
__device__ inline double intra_pair_term(const DevConst& dc, double qa, double qb, const double* pa, const double* pb) {
    double dx = pb[0] - pa[0], dy = pb[1] - pa[1], dz = pb[2] - pa[2];
    if (dc.tri) min_image_tri_brick(dc, dx, dy, dz);   // sites of one molecule are closer than half the shortest cell edge:
    else {                                             // the brick image is the nearest one, no 27-image search (and no call here)
        dx = min_image(dx, dc.L[0], dc.invL[0]); dy = min_image(dy, dc.L[1], dc.invL[1]); dz = min_image(dz, dc.L[2], dc.invL[2]);
    }
    const double r = sqrt(dx * dx + dy * dy + dz * dz);
    return r > MMC_ERR ? qa * qb * (erfc(dc.alpha * r) - 1.0) / r : 0.0;
}
// pre != nullptr: the pair terms were formed in parallel while the molecule was staged (stage_molecule), in the order of the
// loops below; the sum is the same, bit for bit
__device__ inline double intra_energy(const DevConst& dc, int nat, const double* q, const double (*p)[3], const double* pre = nullptr) {
    double e = 0.0;
    int k = 0;
    for (int a1 = 0; a1 < nat - 1; a1++)
        for (int a2 = a1 + 1; a2 < nat; a2++, k++)
            e += pre ? pre[k] : intra_pair_term(dc, q[a1], q[a2], p[a1], p[a2]);
    return e * dc.coul_const;
}

// Assembles type(energy_state) old/new from the six sweep sums exactly as ComputeOldEnergy /
// ComputeNewEnergy do for each move type (monte_carlo_utils.f90:275-396) and the acceptance probability.
// v = {lj_old, coul_old, lj_new, coul_new, sum fW|A|^2, sum fW|A'|^2}.  Single thread.
__device__ inline void finalize_trial(const DevConst& dc, const ChainState& cs, int c, Trial& t, const double* v, const double* intra_pre = nullptr) {
    // intra_pre: pair terms of the created / deleted molecule's intra-residue sum, formed by stage_molecule (batched kernels)
    if (t.kind <= 0) return;
    const double lj_o = v[0], c_o = v[1] * dc.coul_const, lj_n = v[2], c_n = v[3] * dc.coul_const;
    const double rec_o = v[4] * dc.recip_const, rec_n = v[5] * dc.recip_const;
    double* o = t.old_e; double* n = t.new_e;
    const double e_recip_tracked = cs.energy[c * MMC_NE + MMC_E_RECIP];
    if (t.kind == MMC_CREATION) {
        o[0] = 0; o[1] = 0; o[3] = 0; o[4] = 0; o[2] = e_recip_tracked;
        o[5] = o[0] + o[1] + o[2] + o[3] + o[4];
        n[2] = rec_n; n[0] = lj_n; n[1] = c_n; n[3] = dc.self_e[t.res]; n[4] = intra_energy(dc, t.nat, t.q, t.p_new, intra_pre);
        n[5] = n[0] + n[1] + n[2] + n[3] + n[4];
    } else if (t.kind == MMC_SWAP) {
        o[3] = dc.self_e[t.res]; o[4] = intra_energy(dc, t.nat, t.q, t.p_old); o[0] = lj_o; o[1] = c_o; o[2] = e_recip_tracked;
        o[5] = o[0] + o[1] + o[2] + o[3] + o[4];
        n[2] = rec_n; n[0] = lj_n; n[1] = c_n; n[3] = dc.self_e[t.res2]; n[4] = intra_energy(dc, t.nat2, t.q2, t.p_new);
        n[5] = n[0] + n[1] + n[2] + n[3] + n[4];
    } else if (t.kind == MMC_DELETION) {
        o[3] = dc.self_e[t.res]; o[4] = intra_energy(dc, t.nat, t.q, t.p_old, intra_pre); o[0] = lj_o; o[1] = c_o; o[2] = e_recip_tracked;
        o[5] = o[0] + o[1] + o[2] + o[3] + o[4];
        n[0] = 0; n[1] = 0; n[3] = 0; n[4] = 0; n[2] = rec_n;
        n[5] = n[0] + n[1] + n[2] + n[3] + n[4];
    } else {
        o[2] = rec_o; o[0] = lj_o; o[1] = c_o; o[3] = 0; o[4] = 0; o[5] = o[0] + o[1] + o[2];
        n[2] = rec_n; n[0] = lj_n; n[1] = c_n; n[3] = 0; n[4] = 0; n[5] = n[0] + n[1] + n[2];
    }
    // mc_acceptance_probability monte_carlo_utils.f90:184-221; N is read after the count update
    const int cnt = cs.count[c * MMC_MAX_RES + t.res];
    const double V = dc.volume, phi = cs.fug[c * MMC_MAX_RES + t.res], T = cs.temp[c];
    const double de = n[5] - o[5];
    if (t.kind == MMC_SWAP) {  // mc_acceptance_probability_swap monte_carlo_utils.f90:228-268, counts before the move
        const double phi2 = cs.fug[c * MMC_MAX_RES + t.res2], n2 = (double)cs.count[c * MMC_MAX_RES + t.res2];
        t.prob = fmin(1.0, (phi2 / phi) * ((double)cnt / (n2 + 1.0)) * exp(-de / T));
    } else
    if (t.kind == MMC_CREATION) t.prob = fmin(1.0, (phi * V / (double)(cnt + 1)) * exp(-de / T));
    else if (t.kind == MMC_DELETION) t.prob = fmin(1.0, (((double)(cnt - 1) + 1.0) / (phi * V)) * exp(-de / T));
    else t.prob = fmin(1.0, exp(-de / T));
}

// AcceptMove / AcceptCreationMove / AcceptDeletionMove (monte_carlo_utils.f90:410, create_molecule.f90:96,
// delete_molecule.f90:126).  Rejection needs no work: coordinates are only written here and A(k) is
// double buffered.  Single thread.
__device__ inline void commit_trial(const DevConst& dc, const ChainState& cs, int c, Trial& t, int accept) {
    if (t.kind <= 0) return;
    t.accepted = accept;
    const double2 nan2 = make_double2(__longlong_as_double(-1LL), __longlong_as_double(-1LL));
    double2* fwc = cs.fw_pair + (size_t)c * dc.g_mols;
    if (!accept) {
        // a rejected move leaves the geometry: remember the old geometry's framework sums if this trial had to form them
        if (dc.fw_cache && t.fw_old_fresh && t.kind != MMC_CREATION && t.kind != MMC_SWAP)
            fwc[dc.g_mol_base[t.res] + t.mol] = make_double2(t.fw_old[0], t.fw_old[1]);
        t.kind = -t.kind; return;  // negative kind = finished trial (cannot be committed twice)
    }
    double* E = cs.energy + c * MMC_NE;
    const double* o = t.old_e; const double* n = t.new_e;
    mmc_counters& ctr = cs.counters[c];
    const int r = t.res, m = t.mol, nat = t.nat;
    int& cnt = cs.count[c * MMC_MAX_RES + r];
    if (t.kind == MMC_TRANSLATION || t.kind == MMC_ROTATION) {
        E[MMC_E_RECIP] = E[MMC_E_RECIP] + n[2] - o[2];
        E[MMC_E_NONCOUL] = E[MMC_E_NONCOUL] + n[0] - o[0];
        E[MMC_E_COUL] = E[MMC_E_COUL] + n[1] - o[1];
        E[MMC_E_TOTAL] = E[MMC_E_TOTAL] + n[5] - o[5];
        if (t.kind == MMC_TRANSLATION) ctr.translations++; else ctr.rotations++;
        if (dc.fw_cache) fwc[dc.g_mol_base[r] + m] = t.fw_new_ok ? make_double2(t.fw_new[0], t.fw_new[1]) : nan2;   // the new geometry's framework sums
        double* com = g_com(dc, cs, c, r, m);
        for (int d = 0; d < 3; d++) com[d] = t.com_new[d];
        for (int a = 0; a < nat; a++) {
            const size_t ia = g_atom(dc, c, r, m, a);
            for (int d = 0; d < 3; d++) cs.goff[ia * 3 + d] = t.off_new[a][d];
            cs.gx[ia] = t.p_new[a][0]; cs.gy[ia] = t.p_new[a][1]; cs.gz[ia] = t.p_new[a][2];
        }
    } else {
        E[MMC_E_RECIP] = n[2];
        E[MMC_E_NONCOUL] = E[MMC_E_NONCOUL] + n[0] - o[0];
        E[MMC_E_COUL] = E[MMC_E_COUL] + n[1] - o[1];
        E[MMC_E_SELF] = E[MMC_E_SELF] + n[3] - o[3];
        E[MMC_E_INTRA] = E[MMC_E_INTRA] + n[4] - o[4];
        E[MMC_E_TOTAL] = E[MMC_E_TOTAL] + n[5] - o[5];
        if (t.kind == MMC_CREATION) {
            ctr.creations++; cnt = cnt + 1;  // geometry already in the slot
            if (dc.fw_cache) fwc[dc.g_mol_base[r] + m] = t.fw_new_ok ? make_double2(t.fw_new[0], t.fw_new[1]) : nan2;
            int& nr = cs.rcount[c * MMC_MAX_RES + r];
            if (dc.has_reservoir && nr > 0 && t.rsv >= 0) {  // (strict mode: the host owns the pool, rsv < 0)  AcceptCreationMove create_molecule.f90:118-130: pool slot <- last pool slot
                double* ro = cs.roff + ((size_t)c * dc.r_atoms + dc.r_atom_base[r]) * 3;
                for (int i = 0; i < nat * 3; i++) ro[(size_t)t.rsv * nat * 3 + i] = ro[(size_t)(nr - 1) * nat * 3 + i];
                nr = nr - 1;
            }
        } else {
            if (t.kind == MMC_SWAP) {   // inserted geometry already in its slot
                if (dc.fw_cache) fwc[dc.g_mol_base[t.res2] + cs.count[c * MMC_MAX_RES + t.res2]] = nan2;
                ctr.swaps++; cs.count[c * MMC_MAX_RES + t.res2] += 1;
            }
            else ctr.deletions++;
            const int last = cnt - 1;  // RemoveMolecule delete_molecule.f90:99-118: slot[m] <- slot[last]
            if (t.kind == MMC_DELETION && dc.has_reservoir) {
                // AcceptDeletionMove delete_molecule.f90:148-166 appends the offsets of slot last_molecule_index (not of the
                // deleted slot) to the pool; its three position draws are consumed by the caller
                int& nr = cs.rcount[c * MMC_MAX_RES + r];
                if (nr < dc.r_cap[r]) {
                    double* ro = cs.roff + ((size_t)c * dc.r_atoms + dc.r_atom_base[r] + (size_t)nr * nat) * 3;
                    for (int a = 0; a < nat; a++) for (int d = 0; d < 3; d++) ro[a * 3 + d] = cs.goff[g_atom(dc, c, r, last, a) * 3 + d];
                    nr = nr + 1;
                }
            }
            if (m != last) {
                if (dc.fw_cache) fwc[dc.g_mol_base[r] + m] = fwc[dc.g_mol_base[r] + last];   // the cache entry moves with the molecule
                double* com = g_com(dc, cs, c, r, m); const double* cl = g_com(dc, cs, c, r, last);
                for (int d = 0; d < 3; d++) com[d] = cl[d];
                for (int a = 0; a < nat; a++) {
                    const size_t ia = g_atom(dc, c, r, m, a), il = g_atom(dc, c, r, last, a);
                    for (int d = 0; d < 3; d++) cs.goff[ia * 3 + d] = cs.goff[il * 3 + d];
                    cs.gx[ia] = cs.gx[il]; cs.gy[ia] = cs.gy[il]; cs.gz[ia] = cs.gz[il];
                }
            }
            cnt = cnt - 1;
        }
    }
    cs.cur[c] ^= 1;  // A'(k) becomes A(k)
    t.kind = -t.kind;
}

// AdjustMoveStepSizes monte_carlo_utils.f90:99-129, once per block for every chain (batched mode).  Literal: the
// acceptance ratios are formed in single precision (real(counter)), `else if (acc - target < tol)` also fires inside the
// tolerance band, and the rotation branch is min(angle * 1.95, MIN_ROTATION_ANGLE) (quirks Q6).
__global__ void k_adjust_step_sizes(DevConst dc, ChainState cs) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= dc.n_chains) return;
    const mmc_counters& k = cs.counters[c];
    const double target = 0.40, tol = 0.05;          // parameters.f90:24-25
    if (k.trial_translations > 500) {                 // MIN_TRIALS_FOR_RECALIBRATION parameters.f90:30
        const double acc = (double)((float)k.translations / (float)k.trial_translations);
        if (acc - target > tol) cs.tstep[c] = fmin(cs.tstep[c] * 1.05, 3.0);         // MAX_TRANSLATION_STEP
        else if (acc - target < tol) cs.tstep[c] = fmax(cs.tstep[c] * 0.95, 1.0e-3);  // MIN_TRANSLATION_STEP
    }
    if (k.trial_rotations > 500) {
        const double acc = (double)((float)k.rotations / (float)k.trial_rotations);
        if (acc - target > tol) cs.rstep[c] = fmin(cs.rstep[c] * 1.05, 0.78);         // MAX_ROTATION_ANGLE
        else if (acc - target < tol) cs.rstep[c] = fmin(cs.rstep[c] * 1.95, 1.0e-3);  // as written in the reference
    }
}

struct BatchCtl { RngSrc rng; mmc_move_record* recs; long long rec_stride; long long rec_base; };

// strict mode: sums the per-CTA partials of k_sweep in a fixed order -> Trial.old_e/new_e ; the host decides
__global__ void __launch_bounds__(TAIL_THREADS) k_finalize(DevConst dc, ChainState cs, int chain, double* out /* 13 doubles */) {
    __shared__ double red[6 * TAIL_THREADS / 32];
    double v[6] = {0, 0, 0, 0, 0, 0};
    const double* pp = cs.part + (size_t)chain * dc.sweep_parts * 6;
    for (int i = threadIdx.x; i < dc.sweep_parts; i += TAIL_THREADS)
        for (int k = 0; k < 6; k++) v[k] += pp[i * 6 + k];
    block_sum<6, TAIL_THREADS>(v, red);
    if (threadIdx.x == 0) {
        finalize_trial(dc, cs, chain, cs.trial[chain], v);
        if (out) {
            const Trial& t = cs.trial[chain];
            for (int i = 0; i < MMC_NE; i++) { out[i] = t.old_e[i]; out[MMC_NE + i] = t.new_e[i]; }
            out[2 * MMC_NE] = t.prob;
        }
    }
}
__global__ void k_commit(DevConst dc, ChainState cs, int chain, int accept) {
    if (threadIdx.x == 0) commit_trial(dc, cs, chain, cs.trial[chain], accept);
}

// ---- k-sharded mode: ONE system (or a few chains) spread over several GPUs (optional mode) ------------
// Every rank holds the coordinates and makes the SAME proposal from the same Philox counters; rank r sweeps 1/G of the
// framework tiles and owns 1/G of the (kx, ky) rows of A(k) (k_sweep with dc.shard_rank / dc.shard_count).  The six
// partial sums of every chain's trial are all-reduced by the caller (NCCL over NVLink, 6 x chains doubles per MC
// step); every rank then takes the identical accept decision, so no broadcast follows.
//   k_propose_all : grid = chains, proposal of this step
//   k_part_sum    : grid = chains, this rank's sweep partials -> sums[chain][6] (fixed order)
//   k_finish      : grid = chains, energy assembly from the REDUCED sums, accept draw, commit
__global__ void k_propose_all(DevConst dc, ChainState cs, const BatchCtl* ctl, int c0) {
    if (threadIdx.x != 0) return;
    const int c = c0 + blockIdx.x;
    RngSrc rs = ctl->rng;
    propose_move(dc, cs, rs, c, cs.trial[c]);
    if (cs.trial[c].error) atomicOr(cs.error_flag, MMC_ERRF_CAPACITY);
}
__global__ void __launch_bounds__(TAIL_THREADS) k_part_sum(DevConst dc, ChainState cs, int c0, double* sums /* [chains][6] */) {
    __shared__ double red[6 * TAIL_THREADS / 32];
    const int c = c0 + blockIdx.x;
    double v[6] = {0, 0, 0, 0, 0, 0};
    const double* pp = cs.part + (size_t)c * dc.sweep_parts * 6;
    if (cs.trial[c].kind > 0)
        for (int i = threadIdx.x; i < dc.sweep_parts; i += TAIL_THREADS)
            for (int k = 0; k < 6; k++) v[k] += pp[i * 6 + k];
    block_sum<6, TAIL_THREADS>(v, red);
    if (threadIdx.x == 0) for (int k = 0; k < 6; k++) sums[(size_t)blockIdx.x * 6 + k] = v[k];
}
// decide != 0: draw the accept uniform after the proposal's draws and commit (batched sharded step);
// decide == 0: only assemble old/new energies into out (strict sharded trial, the host decides)
__global__ void k_finish(DevConst dc, ChainState cs, const BatchCtl* ctl, int c0, const double* sums, int decide, double* out) {
    if (threadIdx.x != 0) return;
    const int c = c0 + blockIdx.x;
    Trial& t = cs.trial[c];
    const int kind = t.kind;
    if (kind > 0) {
        finalize_trial(dc, cs, c, t, sums + (size_t)blockIdx.x * 6);
        if (decide) {
            RngSrc rs = ctl->rng;
            Rng rng{&rs, cs, c, t.draw};
            const double u = rng.next();
            const int acc = u <= t.prob ? 1 : 0;
            if (acc && kind == MMC_DELETION && dc.has_reservoir) { rng.next(); rng.next(); rng.next(); }  // reservoir placement draws
            commit_trial(dc, cs, c, t, acc);
        }
    }
    if (out) {
        for (int i = 0; i < MMC_NE; i++) { out[i] = t.old_e[i]; out[MMC_NE + i] = t.new_e[i]; }
        out[2 * MMC_NE] = t.prob;
    }
    if (decide && ctl->recs) {
        mmc_move_record& rec = ctl->recs[(size_t)c * ctl->rec_stride + ctl->rec_base];
        rec.kind = kind > 0 ? kind : 0; rec.res = kind > 0 ? (kind == MMC_SWAP ? t.res + 256 * t.res2 : t.res) : 0; rec.mol = kind > 0 ? t.mol : 0;
        rec.accepted = kind > 0 ? t.accepted : 0; rec.prob = kind > 0 ? t.prob : 0.0;
        for (int i = 0; i < MMC_NE; i++) { rec.old_e[i] = kind > 0 ? t.old_e[i] : 0.0; rec.new_e[i] = kind > 0 ? t.new_e[i] : 0.0; }
    }
}


// positions = COM + offset for every guest slot of every chain (after a bulk upload); grid = (tiles, n_chains)
__global__ void k_rebuild_positions(DevConst dc, ChainState cs) {
    const int c = blockIdx.y, g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= dc.g_atoms) return;
    int r = 0;
    for (int i = 0; i < dc.n_active; i++) { const int rr = dc.active_list[i]; if (g >= dc.g_atom_base[rr]) r = rr; }
    const int m = (g - dc.g_atom_base[r]) / dc.natoms[r];
    const size_t ia = (size_t)c * dc.g_atoms + g;
    const double* com = cs.gcom + ((size_t)c * dc.g_mols + dc.g_mol_base[r] + m) * 3;
    const double* off = cs.goff + ia * 3;
    const double x = com[0] + off[0], y = com[1] + off[1], z = com[2] + off[2];
    // framework-energy cache: an entry stays valid exactly when every atom of its molecule is, bit for bit, where it was (the
    // same sum com + off as build_trial's p_new): a host that uploads what it downloaded keeps the cache
    if (dc.fw_cache && (x != cs.gx[ia] || y != cs.gy[ia] || z != cs.gz[ia]))
        cs.fw_pair[(size_t)c * dc.g_mols + dc.g_mol_base[r] + m] = make_double2(__longlong_as_double(-1LL), __longlong_as_double(-1LL));
    cs.gx[ia] = x; cs.gy[ia] = y; cs.gz[ia] = z;
}
__global__ void k_advance(BatchCtl* ctl, int g) { ctl->rng.step += (unsigned long long)g; ctl->rec_base += g; }
