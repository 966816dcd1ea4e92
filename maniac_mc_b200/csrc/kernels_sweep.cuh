// kernels_sweep.cuh -- the two sweeps of one trial move, written once and used by both drivers:
//   * strict mode  : k_sweep, dc.sweep_parts CTAs share one chain's trial (whole GPU on one move)
//   * batched mode : k_chain_run (kernels_chain.cuh), one CTA owns a chain and loops over MC steps
//
// Real-space sweep = ComputePairInteractionEnergy_singlemol (energy_utils.f90:374-442) for the old and
// the proposed geometry in ONE pass.  The framework is stored in Morton order in tiles of 32 atoms with a
// bounding box each; a tile is visited only if it can hold an atom within `reach` of the moved molecule
// (reach = max(r_c, x_cut/alpha): LJ needs r < r_c, and erfc(alpha r) < erfc(x_cut) = 2e-17 (x_cut = 6) beyond it, which is
// below one ulp of the sums the reference forms over ALL minimum-image pairs, Q3).  With culling disabled
// (reach2 = inf) every pair is evaluated exactly as the reference does.
//
// Reciprocal sweep = ComputeRecipEnergySingleMol (ewald_energy.f90:191-274), parallel over k:
// A'(k) = A(k) + sum_terms X_t(kx) Y_t(ky) qZ_t(kz), terms = atoms added (+q) and atoms removed (-q);
// the 1-D tables (ComputePhaseFactors1D, ewald_phase.f90:90-109) live in shared memory, extended to negative
// k by conjugation as the reference does (:105-107).
#pragma once
#include "common.cuh"
#include "kernels_move.cuh"

// -DMMC_PHASE_TIMING: thread 0 of every CTA timestamps the phase boundaries of each trial (clock64) and the totals are added
// to cs.phase_cycles at the end of the launch (mmc_dbg_phase_cycles).  Diagnostic builds only: the stamps cost registers.
// slots: 0 propose, 1 stage (terms + molecule), 2 cull + real-space sweep, 3 block sum, 4 reciprocal sweep, 5 block sum,
//        6 finalize + accept + commit + record, 7 trials
#ifdef MMC_PHASE_TIMING
#define MMC_TSTAMP(slot) do { if (threadIdx.x == 0) { const long long t_ = clock64(); s.mol->phase[slot] += (unsigned long long)(t_ - s.mol->t_last); s.mol->t_last = t_; } } while (0)
#else
#define MMC_TSTAMP(slot) do { } while (0)
#endif

// moved molecule staged in shared memory
struct MolS {
    double po[MMC_MAX_MOL_ATOMS][3], pn[MMC_MAX_MOL_ATOMS][3];
    double q[MMC_MAX_MOL_ATOMS];
    double cen[2][4];   // bounding sphere {x, y, z, radius + reach} of the old / new geometry
    double cc[4];       // common sphere of the geometries swept: {x, y, z, (radius + reach)^2}
    double ccf[4];      // the same for the FRAMEWORK sweep (new geometry only when the old one's framework energy is cached)
    double fw_old[2];   // framework part {LJ, Coulomb (reduced)} of the old geometry: from the cache, or summed in this trial
    // pair terms of the intra-residue sum of a created / deleted molecule (intra_ok), formed by threads 96.. while the molecule is
    // staged instead of by thread 0 in finalize_trial (6 erfc + sqrt + division sequences for TIP4P: r2zk, ~20 k cycles of the
    // single-thread section of every insertion / deletion)
    double intra[MMC_MAX_MOL_ATOMS * (MMC_MAX_MOL_ATOMS - 1) / 2];
    int intra_ok;
    double fws[4];      // framework-energy cache on (cache_on): this CTA's FRAMEWORK partial sums {lj_old, coul_old, lj_new, coul_new};
                        // sums[0..3] then hold the guests' part only (fold_fw_cache puts them together)
    int cache_on;
    int shift_okf;      // shift_ok of the sphere ccf
    int fo;             // framework sweep evaluates the old geometry (do_old and not cached)
    int fw_cached;      // fw_old came from the cache
    int shift_ok;       // reach + radius < L/2 on every axis: one periodic shift per other atom is exact
    int type[MMC_MAX_MOL_ATOMS];
    int nat, kind, do_old, do_new, excl, res, mol, nterm, nlist, cur;
    int fact;           // MMC_FACTORED_TRANSLATION: terms = +q phi(old) of the charged atoms, last term = e^{i k d}
    signed char term[2 * MMC_MAX_MOL_ATOMS];  // reciprocal terms: +(a+1) = atom a added, -(a+1) = atom a removed
    double sums[6];        // block sums of this trial: parked here between the phases so they hold no registers
    unsigned int work[2];  // executed real-space work of this CTA (distances formed, pairs inside reach)
    double u[16];          // this step's Philox uniforms, drawn ahead by 16 lanes (MMC_PRE_DRAWS)
#ifdef MMC_PHASE_TIMING
    long long t_last; unsigned long long phase[8];
#endif
};

struct SweepCtx {
    double* gtab;    // [MMC_GTAB_DEG+1][MMC_GTAB_MAXSEG] Coulomb kernel table
    double2* lj;     // [max_nat][ntypes] {sigma^2, 4 eps} rows of the moved molecule's atom types
    double2 *X, *Y;  // [2 max_nat][nx | ny] 1-D phase tables, one row per reciprocal term
#ifdef MMC_RECIP_DMMA
    // q_t Z_t(kz), kz >= 0 only, as two real planes [tpad][zS] (term-major): the A operand of the FP64 MMA of the
    // reciprocal sweep.  tpad = terms padded to a multiple of 4, zS = row stride (kz padded to a multiple of 8, + 4: bank
    // spread of the fragment load, recip_zstride); the pads that are read are zero.
    double *ZR, *ZI;
    int tpad, zS, nzb;    // nzb = blocks of 8 kz
#else
    double2* Z;      // [kmax_z + 1][zs] q_t Z_t(kz), kz >= 0 only, term fastest; zs = 2 max_nat + 1 (odd: no bank conflicts)
    double2* XY;     // [2 buffers][threads / 4 rows][zs] products X_t(kx) Y_t(ky) of a chunk of (kx, ky) rows, term fastest
#endif
    uint32_t* mask;  // [CULL_CHUNK / 32]
    int* prefix;     // [CULL_CHUNK / 32]
    uint16_t* list;  // [CULL_CHUNK]
    double* red;     // [8 * warps]
    MolS* mol;
    Trial* trial;         // the pending trial of this CTA's chain (batched kernel, fused strict-mode kernel)
    unsigned char* cscr;  // [warps][32] lane ids exchanged by compact_push
    int nx, ny, nz, zs;   // nz = kmax_z + 1
};

__host__ __device__ inline int recip_tpad(int max_nat) { return ((2 * max_nat + 3) / 4) * 4; }
__host__ __device__ inline int recip_zstride(int kmax_z) {
    // kz padded to a multiple of 8, plus 4: stride = 4 or 12 (mod 16) doubles, so that the 16 lanes of a half warp
    // (4 terms x 4 kz, an 8-byte load is served per half warp) fall into 16 distinct 8-byte bank pairs
    return ((kmax_z + 1 + 7) / 8) * 8 + 4;
}
__host__ __device__ inline size_t sweep_smem_layout(int max_nat, int ntypes, const int kmax[3], int threads, size_t off[12]) {
    size_t o = 0;
    const int nx = kmax[0] + 1, ny = 2 * kmax[1] + 1, nz = kmax[2] + 1, zs = 2 * max_nat + 1;
    off[0] = o; o += (size_t)(MMC_GTAB_DEG + 1) * MMC_GTAB_MAXSEG * sizeof(double);
    off[1] = o; o += (size_t)max_nat * ntypes * sizeof(double2);
    off[2] = o; o += (size_t)2 * max_nat * nx * sizeof(double2);
    off[3] = o; o += (size_t)2 * max_nat * ny * sizeof(double2);
#ifdef MMC_RECIP_DMMA
    off[4] = o; o += (size_t)2 * recip_tpad(max_nat) * recip_zstride(kmax[2]) * sizeof(double);
    (void)nz; (void)zs;
#else
    off[4] = o; o += (size_t)nz * zs * sizeof(double2);
#endif
    off[5] = o; o += (size_t)8 * (threads / 32) * sizeof(double);
    off[6] = o; o += (sizeof(MolS) + 15) / 16 * 16;
    off[7] = o; o += (CULL_CHUNK / 32) * sizeof(uint32_t) * 2;
    off[8] = o; o += CULL_CHUNK * sizeof(uint16_t);
    o = (o + 15) / 16 * 16;
#ifdef MMC_RECIP_DMMA
    off[9] = o; (void)threads;     // no staging buffer: the (kx, ky) products are formed in registers
#else
    off[9] = o; o += (size_t)2 * (threads / 4) * zs * sizeof(double2);
#endif
    off[10] = o; o += (sizeof(Trial) + 15) / 16 * 16;
    off[11] = o; o += (size_t)(threads / 32) * 32;   // compacted sweep: 32 lane ids per warp
    return o;
}

__device__ __forceinline__ SweepCtx sweep_carve(const DevConst& dc, unsigned char* smem, int threads) {
    size_t off[12];
    sweep_smem_layout(dc.max_nat, dc.ntypes, dc.kmax, threads, off);
    SweepCtx s;
    s.gtab = (double*)(smem + off[0]); s.lj = (double2*)(smem + off[1]);
    s.X = (double2*)(smem + off[2]); s.Y = (double2*)(smem + off[3]);
#ifdef MMC_RECIP_DMMA
    s.tpad = recip_tpad(dc.max_nat); s.zS = recip_zstride(dc.kmax[2]); s.nzb = (dc.kmax[2] + 8) >> 3;
    s.ZR = (double*)(smem + off[4]); s.ZI = s.ZR + (size_t)s.tpad * s.zS;
#else
    s.Z = (double2*)(smem + off[4]);
#endif
    s.red = (double*)(smem + off[5]); s.mol = (MolS*)(smem + off[6]); s.trial = (Trial*)(smem + off[10]); s.cscr = smem + off[11];
    s.mask = (uint32_t*)(smem + off[7]); s.prefix = (int*)(s.mask + CULL_CHUNK / 32);
    s.list = (uint16_t*)(smem + off[8]);
#ifndef MMC_RECIP_DMMA
    s.XY = (double2*)(smem + off[9]);
#endif
    s.nx = dc.kmax[0] + 1; s.ny = 2 * dc.kmax[1] + 1; s.nz = dc.kmax[2] + 1; s.zs = 2 * dc.max_nat + 1;
    return s;
}

// once per kernel: the Coulomb kernel table
__device__ __forceinline__ void sweep_load_tables(const DevConst& dc, const SweepCtx& s) {
    for (int i = threadIdx.x; i < (MMC_GTAB_DEG + 1) * MMC_GTAB_MAXSEG; i += blockDim.x) s.gtab[i] = dc.gtab[i];
}

// Stages the moved molecule of the pending Trial of chain c for the real-space sweep: positions, charges, LJ rows,
// bounding spheres.  The Trial carries an OLD atom set (nat, q, type, p_old) and a NEW atom set (nat2, q2, type2,
// p_new); they are the same molecule for translation / rotation / insertion / deletion and differ for a swap.
//   pass 0: both sets in one sweep (every move but the swap)
//   pass 1: swap, removed molecule only (old geometry)      pass 2: swap, inserted molecule only (new geometry)
// All threads; ends with a barrier.
__device__ __forceinline__ void stage_molecule(const DevConst& dc, const ChainState& cs, int c, const SweepCtx& s, const Trial& t, int pass, bool use_cache = false) {
    MolS& m = *s.mol;
    const int tid = threadIdx.x;
    const bool use_new_set = pass == 2;
    const int nat = use_new_set ? t.nat2 : t.nat;
    const double* q = use_new_set ? t.q2 : t.q;
    const int* type = use_new_set ? t.type2 : t.type;
    const bool uo = pass == 0 ? t.do_old != 0 : pass == 1, un = pass == 0 ? t.do_new != 0 : pass == 2;
    if (tid == 0) {
        m.nat = nat; m.kind = t.kind; m.do_old = uo; m.do_new = un; m.excl = t.excl; m.res = t.res; m.mol = t.mol;
        m.cur = cs.cur[c];
        const bool cached = use_cache && pass == 0 && uo && t.fw_old_ok;
        m.fw_cached = cached; m.fo = uo && !cached;
        m.fw_old[0] = cached ? t.fw_old[0] : 0.0; m.fw_old[1] = cached ? t.fw_old[1] : 0.0;
    }
    if (tid == 0) m.intra_ok = pass == 0 && (t.kind == MMC_CREATION || t.kind == MMC_DELETION);
    if (t.kind <= 0) { __syncthreads(); return; }
    if (pass == 0 && (t.kind == MMC_CREATION || t.kind == MMC_DELETION) && tid >= 96 && tid - 96 < nat * (nat - 1) / 2) {
        int k = tid - 96, a1 = 0;
        while (k >= nat - 1 - a1) { k -= nat - 1 - a1; a1++; }     // pair index -> (a1, a2) in intra_energy's loop order
        const int a2 = a1 + 1 + k;
        const double (*p)[3] = t.kind == MMC_CREATION ? t.p_new : t.p_old;
        m.intra[tid - 96] = intra_pair_term(dc, q[a1], q[a2], p[a1], p[a2]);
    }
    if (tid < nat) {
        for (int d = 0; d < 3; d++) { m.po[tid][d] = t.p_old[tid][d]; m.pn[tid][d] = t.p_new[tid][d]; }
        m.q[tid] = q[tid]; m.type[tid] = type[tid];
    }
    if (tid < 2) {  // bounding sphere of one geometry: centroid + largest distance + reach
        const double (*p)[3] = tid == 0 ? t.p_old : t.p_new;
        double cx = 0, cy = 0, cz = 0;
        for (int a = 0; a < nat; a++) { cx += p[a][0]; cy += p[a][1]; cz += p[a][2]; }
        cx /= nat; cy /= nat; cz /= nat;
        double r2 = 0;
        for (int a = 0; a < nat; a++) {
            const double dx = p[a][0] - cx, dy = p[a][1] - cy, dz = p[a][2] - cz;
            r2 = fmax(r2, dx * dx + dy * dy + dz * dz);
        }
        m.cen[tid][0] = cx; m.cen[tid][1] = cy; m.cen[tid][2] = cz; m.cen[tid][3] = sqrt(r2) * (1.0 + 1e-12) + dc.reach;
    }
    if (tid == 32 || tid == 64) {  // common sphere of the swept geometries (32: all atoms; 64, another warp: the framework sweep,
        // which skips a cached old geometry), and whether the single-shift path is exact
        const bool uof = tid == 32 ? uo : (uo && !(use_cache && pass == 0 && t.fw_old_ok));
        double cx = 0, cy = 0, cz = 0; int n = 0;
        for (int a = 0; a < nat; a++) {
            if (uof) { cx += t.p_old[a][0]; cy += t.p_old[a][1]; cz += t.p_old[a][2]; n++; }
            if (un) { cx += t.p_new[a][0]; cy += t.p_new[a][1]; cz += t.p_new[a][2]; n++; }
        }
        if (n) { cx /= n; cy /= n; cz /= n; }
        double r2 = 0;
        for (int a = 0; a < nat; a++) {
            if (uof) { const double dx = t.p_old[a][0] - cx, dy = t.p_old[a][1] - cy, dz = t.p_old[a][2] - cz; r2 = fmax(r2, dx * dx + dy * dy + dz * dz); }
            if (un) { const double dx = t.p_new[a][0] - cx, dy = t.p_new[a][1] - cy, dz = t.p_new[a][2] - cz; r2 = fmax(r2, dx * dx + dy * dy + dz * dz); }
        }
        const double R = sqrt(r2) * (1.0 + 1e-12) + 1e-9 + dc.reach;
        double* cc = tid == 32 ? m.cc : m.ccf;
        cc[0] = cx; cc[1] = cy; cc[2] = cz; cc[3] = R * R;
        (tid == 32 ? m.shift_ok : m.shift_okf) = R < 0.4999 * fmin(dc.L[0], fmin(dc.L[1], dc.L[2]));
    }
    for (int i = tid; i < nat * dc.ntypes; i += blockDim.x) {
        const int a = i / dc.ntypes, tj = i % dc.ntypes;
        const double2 v = dc.lj[type[a] * dc.ntypes + tj];
        s.lj[i] = make_double2(v.x * v.x, v.y);
    }
    __syncthreads();
}

// Stages the reciprocal terms of the pending Trial: the term map and the 1-D phase tables (ComputeAtomPhase +
// ComputePhaseFactors1D, ewald_phase.f90:41-109: direct cos/sin(k theta)), one row per term with the charge and the
// sign folded into Z.  Added atoms come from the NEW set (p_plus, q2), removed atoms from the OLD set (p_old, q).
// All threads; ends with a barrier.
template <bool TRI>
__device__ __forceinline__ void stage_terms(const DevConst& dc, const ChainState& cs, int c, const SweepCtx& s, const Trial& t) {
    MolS& m = *s.mol;
    const int tid = threadIdx.x;
    if (tid == 0) {
        // atoms with q == 0 add exactly zero to A(k) (ewald_energy.f90:241-256 multiplies by the charge): no term
        int n = 0;
        m.fact = 0;
#ifdef MMC_FACTORED_TRANSLATION
        // a translation moves every atom by the same d: sum q phi(new) - sum q phi(old) = S_mol(k) (e^{i k d} - 1) with
        // S_mol = sum q phi(old): n + 1 product terms instead of 2 n
        if (!TRI && t.kind == MMC_TRANSLATION) {
            for (int a = 0; a < t.nat; a++) if (t.q[a] != 0.0) m.term[n++] = (signed char)(-(a + 1));
            if (n > 1) { m.term[n++] = 127; m.fact = 1; } else n = 0;
        }
        if (!m.fact)
#endif
        if (t.kind > 0) {
            if (t.r_plus) for (int a = 0; a < t.nat2; a++) if (t.q2[a] != 0.0) m.term[n++] = (signed char)(a + 1);
            if (t.r_minus) for (int a = 0; a < t.nat; a++) if (t.q[a] != 0.0) m.term[n++] = (signed char)(-(a + 1));
        }
        m.nterm = n;
    }
    __syncthreads();
    if (t.kind <= 0) return;
    const int nterm = m.nterm;
#ifdef MMC_RECIP_DMMA
    // X and Y rows of the nterm terms, then the two Z planes over the PADDED (term, kz) range the MMA fragments read:
    // terms up to the next multiple of 4 and kz up to the next multiple of 8 are written as zeros
    // (at least one K-block of 4 terms is always read, also when no charged atom moves: its pads must be zeros, not stale data)
    const int nxy = s.nx + s.ny, kb4 = max(4, (nterm + 3) & ~3), nzp = s.nzb * 8;
    for (int i = tid; i < nterm * nxy + kb4 * nzp; i += blockDim.x) {
        int term, j, ax, k; double2* dst = nullptr; double *dr = nullptr, *di = nullptr;
        if (i < nterm * nxy) {
            term = i / nxy; j = i - term * nxy;
            if (j < s.nx) { ax = 0; k = j; dst = s.X + term * s.nx + j; }
            else { j -= s.nx; ax = 1; k = j - dc.kmax[1]; dst = s.Y + term * s.ny + j; }
        } else {
            const int u = i - nterm * nxy;
            term = u / nzp; k = u - term * nzp; ax = 2;
            dr = s.ZR + term * s.zS + k; di = s.ZI + term * s.zS + k;
            if (term >= nterm || k >= s.nz) { *dr = 0.0; *di = 0.0; continue; }
        }
        const int ta = m.term[term];
        const bool plus = ta > 0;
        const int a = (plus ? ta : -ta) - 1;
        const double* pp = plus ? t.p_plus[a] : t.p_old[a];
        const double theta = MMC_TWOPI * (TRI ? frac_coord(dc, pp, ax) : dc.invL[ax] * pp[ax]);
        double sn, co; sincos((double)(k < 0 ? -k : k) * theta, &sn, &co);
        if (k < 0) sn = -sn;
        if (ax == 2) { const double q = plus ? t.q2[a] : -t.q[a]; *dr = co * q; *di = sn * q; }
        else *dst = make_double2(co, sn);
    }
#else
    const int per = s.nx + s.ny + s.nz;
    for (int i = tid; i < nterm * per; i += blockDim.x) {
        const int term = i / per; int j = i % per;
        const int ta = m.term[term];
        const bool plus = ta > 0;
        const int a = ta == 127 ? 0 : (plus ? ta : -ta) - 1;
        int ax, k; double2* dst;
        if (j < s.nx) { ax = 0; k = j; dst = s.X + term * s.nx + j; }
        else if (j < s.nx + s.ny) { j -= s.nx; ax = 1; k = j - dc.kmax[1]; dst = s.Y + term * s.ny + j; }
        else { j -= s.nx + s.ny; ax = 2; k = j; dst = s.Z + j * s.zs + term; }
        const double* pp = plus ? t.p_plus[a] : t.p_old[a];
        double theta = MMC_TWOPI * (TRI ? frac_coord(dc, pp, ax) : dc.invL[ax] * pp[ax]);
#ifdef MMC_FACTORED_TRANSLATION
        if (ta == 127) theta = MMC_TWOPI * dc.invL[ax] * (t.p_plus[0][ax] - t.p_old[0][ax]);
#endif
        double sn, co; sincos((double)(k < 0 ? -k : k) * theta, &sn, &co);
        if (k < 0) sn = -sn;
        if (ax == 2) {
            double q = plus ? t.q2[a] : -t.q[a];
#ifdef MMC_FACTORED_TRANSLATION
            if (m.fact) q = ta == 127 ? 1.0 : t.q[a];
#endif
            co *= q; sn *= q;
        }
        *dst = make_double2(co, sn);
    }
#endif
    __syncthreads();
}

// ----------------------------------------------------------------------------------------------
struct PairConst { double Lx, Ly, Lz, iLx, iLy, iLz, reach2, alpha, rc2, smin, smax; int gbase, g1, g2; };

// one (moved atom, other atom) pair at one geometry; SHIFT: the other atom was already moved to the periodic
// image nearest the molecule, so the plain difference IS the minimum image for every pair within reach
template <bool SHIFT, bool TRI>
__device__ __forceinline__ void pair_one(const DevConst& dc, const PairConst& pc, const double* s_gtab, double dx, double dy, double dz,
                                         double qq, bool charged, double2 lj, double& e_lj, double& e_c, unsigned& work) {
    if (!SHIFT) {
        if (TRI) min_image_tri(dc, dx, dy, dz);
        else {
            dx = min_image_fast(dx, pc.Lx, pc.iLx);
            dy = min_image_fast(dy, pc.Ly, pc.iLy);
            dz = min_image_fast(dz, pc.Lz, pc.iLz);
        }
    }
    const double r2 = fma(dx, dx, fma(dy, dy, dz * dz));
    if (r2 < pc.reach2) {
        work += 0x10000u;
        if (charged) {  // CoulombEnergy term erfc(alpha r)/r, energy_utils.f90:425-433
            double g;
            if (r2 >= pc.smin && r2 < pc.smax) g = coul_kernel_tab(r2, s_gtab, pc.gbase, pc.g1, pc.g2);
            else g = coul_kernel_slow(r2, pc.alpha);
            e_c = fma(qq, g, e_c);
        }
        if (r2 < pc.rc2 && lj.y != 0.0) {  // 4 eps [(sigma/r)^12 - (sigma/r)^6], energy_utils.f90:417-423
            const double s2 = lj.x * mmc_rcp(r2);
            const double s6 = s2 * s2 * s2;
            e_lj = fma(lj.y, fma(s6, s6, -s6), e_lj);
        }
    }
}

// the same pair at the OLD and the NEW geometry in one go: every step of the two evaluations is independent of the other
// geometry's, so the two dependent chains (distance -> table index -> 8 FMAs; reciprocal -> sixth power) interleave.
// Per geometry the operations and their order are those of pair_one, the sums go to the same accumulators: bit-identical.
template <bool SHIFT, bool TRI>
__device__ __forceinline__ void pair_two(const DevConst& dc, const PairConst& pc, const double* s_gtab, double dxo, double dyo, double dzo,
                                         double dxn, double dyn, double dzn, double qq, bool charged, double2 lj, double (&acc)[4], unsigned& work) {
    if (!SHIFT) {
        if (TRI) { min_image_tri(dc, dxo, dyo, dzo); min_image_tri(dc, dxn, dyn, dzn); }
        else {
            dxo = min_image_fast(dxo, pc.Lx, pc.iLx); dyo = min_image_fast(dyo, pc.Ly, pc.iLy); dzo = min_image_fast(dzo, pc.Lz, pc.iLz);
            dxn = min_image_fast(dxn, pc.Lx, pc.iLx); dyn = min_image_fast(dyn, pc.Ly, pc.iLy); dzn = min_image_fast(dzn, pc.Lz, pc.iLz);
        }
    }
    const double r2o = fma(dxo, dxo, fma(dyo, dyo, dzo * dzo)), r2n = fma(dxn, dxn, fma(dyn, dyn, dzn * dzn));
    const bool in_o = r2o < pc.reach2, in_n = r2n < pc.reach2;
    if (!(in_o | in_n)) return;
    work += (in_o ? 0x10000u : 0u) + (in_n ? 0x10000u : 0u);
    // a geometry out of reach is evaluated at the other one's distance and discarded (keeps the table index in range)
    const double so = in_o ? r2o : r2n, sn = in_n ? r2n : r2o;
    if (charged) {  // CoulombEnergy term erfc(alpha r)/r, energy_utils.f90:425-433
        double go, gn;
        if (fmin(so, sn) >= pc.smin && fmax(so, sn) < pc.smax) coul_kernel_tab2(so, sn, s_gtab, pc.gbase, pc.g1, pc.g2, go, gn);
        else {
            go = (so >= pc.smin && so < pc.smax) ? coul_kernel_tab(so, s_gtab, pc.gbase, pc.g1, pc.g2) : coul_kernel_slow(so, pc.alpha);
            gn = (sn >= pc.smin && sn < pc.smax) ? coul_kernel_tab(sn, s_gtab, pc.gbase, pc.g1, pc.g2) : coul_kernel_slow(sn, pc.alpha);
        }
        if (in_o) acc[1] = fma(qq, go, acc[1]);
        if (in_n) acc[3] = fma(qq, gn, acc[3]);
    }
    if (lj.y != 0.0) {  // 4 eps [(sigma/r)^12 - (sigma/r)^6], energy_utils.f90:417-423
        const bool lo = in_o && r2o < pc.rc2, ln = in_n && r2n < pc.rc2;
        if (lo | ln) {
            const double s2o = lj.x * mmc_rcp(so), s2n = lj.x * mmc_rcp(sn);
            const double s6o = s2o * s2o * s2o, s6n = s2n * s2n * s2n;
            if (lo) acc[0] = fma(lj.y, fma(s6o, s6o, -s6o), acc[0]);
            if (ln) acc[2] = fma(lj.y, fma(s6n, s6n, -s6n), acc[2]);
        }
    }
}

// one "other" atom against every atom of the moved molecule, old and/or new geometry
// TRI: tilted cell; the OLD / NEW sets are then also masked at run time (one instantiation serves every move type)
// SHIFT path, step 1: move the other atom to the periodic image nearest the molecule's centre (exact for every pair within
// reach, shift_ok) and test it against the molecule's sphere.  Returns false when it is beyond reach of every moved atom.
template <bool TRI>
__device__ __forceinline__ bool pair_shift(const DevConst& dc, const PairConst& pc, const double* cc, double& xj, double& yj, double& zj) {
    double dx = xj - cc[0], dy = yj - cc[1], dz = zj - cc[2];
    if (TRI) {   // the image inside the brick around the molecule's centre
        min_image_tri_brick(dc, dx, dy, dz);
        xj = cc[0] + dx; yj = cc[1] + dy; zj = cc[2] + dz;
    } else {
        const double big = 6755399441055744.0;
        const double nx = fma(dx, pc.iLx, big) - big, ny = fma(dy, pc.iLy, big) - big, nz = fma(dz, pc.iLz, big) - big;
        xj = fma(-nx, pc.Lx, xj); yj = fma(-ny, pc.Ly, yj); zj = fma(-nz, pc.Lz, zj);
        dx = xj - cc[0]; dy = yj - cc[1]; dz = zj - cc[2];
    }
    return !(fma(dx, dx, fma(dy, dy, dz * dz)) > cc[3]);
}

// PREP: the atom was already shifted and sphere-tested by pair_shift (compacted sweep)
template <bool OLD, bool NEW, bool SHIFT, bool TRI, bool PREP = false>
__device__ __forceinline__ void pair_atom(const DevConst& dc, const PairConst& pc, const SweepCtx& s, int ntypes, double xj, double yj, double zj,
                                          double qj, int tj, bool uo, bool un, const double* cc, double (&acc)[4], unsigned& cnt) {
    const MolS& m = *s.mol;
    if (SHIFT && !PREP) {
        if (!pair_shift<TRI>(dc, pc, cc, xj, yj, zj)) return;  // beyond reach of every atom of the molecule
    }
    const bool use_old = OLD && uo, use_new = NEW && un;   // uo / un: run-time masks (compile-time true where the instantiation fixes the set)
    const bool qj_ok = fabs(qj) >= MMC_ERR;
    const int nat = m.nat;
#pragma unroll 2
    for (int a = 0; a < nat; a++) {
        const double qa = m.q[a];
        const bool charged = qj_ok && fabs(qa) >= MMC_ERR;
        const double qq = qa * qj;
        const double2 lj = s.lj[a * ntypes + tj];
        if (!charged && lj.y == 0.0) continue;
        // ComputeDistance: delta = pos(other) - pos(moved)
        cnt += (use_old ? 1u : 0u) + (use_new ? 1u : 0u);
#ifdef MMC_PAIR_TWO   // opt-in: measured 1.7 % SLOWER than the two pair_one calls (r2d A/B, profiles/r2d_ab.log): the compiler already interleaves them
        if (OLD && NEW && !TRI) {   // translation / rotation in an orthorhombic cell: both geometries in lockstep
            pair_two<SHIFT, TRI>(dc, pc, s.gtab, xj - m.po[a][0], yj - m.po[a][1], zj - m.po[a][2], xj - m.pn[a][0], yj - m.pn[a][1], zj - m.pn[a][2],
                                 qq, charged, lj, acc, cnt);
            continue;
        }
#endif
        if (use_old) pair_one<SHIFT, TRI>(dc, pc, s.gtab, xj - m.po[a][0], yj - m.po[a][1], zj - m.po[a][2], qq, charged, lj, acc[0], acc[1], cnt);
        if (use_new) pair_one<SHIFT, TRI>(dc, pc, s.gtab, xj - m.pn[a][0], yj - m.pn[a][1], zj - m.pn[a][2], qq, charged, lj, acc[2], acc[3], cnt);
    }
}

// ---- compacted sweep (SHIFT path), opt-in with -DMMC_COMPACT ----------------------------------------------------------
// Measured NEUTRAL (r2j A/B, profiles/r2j_ab_compact.log: 2.700 vs 2.702 M moves/s): the real-space phase is bound by the BYTES
// its table look-ups move through the shared-memory pipe, and those scale with the active lanes either way; full warps only
// save issue slots, which are not the limiter.  Kept for systems whose tiles are much coarser than the sphere.
// Survivors of the sphere test carried by a warp: lanes [0, n) hold one atom each (already shifted to its nearest image).
struct CompactQ { double x, y, z, q; int t, n; };

// All 32 lanes call this with their candidate atom (`in`: it passed pair_shift).  The survivors are appended to the carried
// set in lane order; every time 32 are held the loop over the moved atoms runs on the full warp.  flush: evaluate whatever
// is left (the last call of a trial).  The lane that holds entry e of the new survivors is found through 32 bytes of
// per-warp shared scratch (each survivor writes its lane id at its rank), the data moves by warp shuffles.
template <bool OLD, bool NEW, bool TRI>
__device__ __forceinline__ void compact_push(const DevConst& dc, const PairConst& pc, const SweepCtx& s, CompactQ& cq, bool in, double xj, double yj,
                                             double zj, double qj, int tj, bool flush, bool uo, bool un, const double* cc, double (&acc)[4], unsigned& cnt) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    unsigned char* scr = s.cscr + (threadIdx.x >> 5) * 32;
    const unsigned mask = __ballot_sync(full, in);
    const int c = __popc(mask);
    if (in) scr[__popc(mask & ((1u << lane) - 1u))] = (unsigned char)lane;
    __syncwarp();
    int n = cq.n, consumed = 0;
    while (consumed < c || (flush && n > 0)) {
        const int take = min(c - consumed, 32 - n);
        if (take > 0) {
            const int e = lane - n;                                   // lanes [n, n + take) take new survivors consumed + e
            const bool mine = e >= 0 && e < take;
            const int src = mine ? (int)scr[consumed + e] : lane;
            const double vx = __shfl_sync(full, xj, src), vy = __shfl_sync(full, yj, src), vz = __shfl_sync(full, zj, src), vq = __shfl_sync(full, qj, src);
            const int vt = __shfl_sync(full, tj, src);
            if (mine) { cq.x = vx; cq.y = vy; cq.z = vz; cq.q = vq; cq.t = vt; }
            n += take; consumed += take;
        }
        if (n == 32 || (flush && consumed == c && n > 0)) {
            if (lane < n) pair_atom<OLD, NEW, true, TRI, true>(dc, pc, s, dc.ntypes, cq.x, cq.y, cq.z, cq.q, cq.t, uo, un, cc, acc, cnt);
            __syncwarp();
            n = 0;
        }
    }
    cq.n = n;
    __syncwarp();   // scr is rewritten by the next call
}

// can tile `t` hold an atom within reach of the moved molecule?  periodic point-to-box distance per axis
template <bool TRI>
__device__ __forceinline__ bool tile_in_reach(const DevConst& dc, const MolS& m, int t, bool use_old, bool use_new) {
    const double* b = dc.tbox + t; const size_t nt = (size_t)dc.n_tiles;
    const double bx = b[0], by = b[nt], bz = b[2 * nt], hx = b[3 * nt], hy = b[4 * nt], hz = b[5 * nt];
    bool hit = false;
    if (TRI) {
        // tilted cell: the z image fixes the y shift and the y image the x shift, so the per-axis bound holds only while
        // the image along each axis is unambiguous (R + h < l/2); otherwise the tile is kept
        for (int g = 0; g < 2; g++) {
            if (g == 0 ? !use_old : !use_new) continue;
            const double R = m.cen[g][3];
            if (R + hx >= 0.5 * dc.L[0] || R + hy >= 0.5 * dc.L[1] || R + hz >= 0.5 * dc.L[2]) return true;
            double dx = m.cen[g][0] - bx, dy = m.cen[g][1] - by, dz = m.cen[g][2] - bz;
            min_image_tri_brick(dc, dx, dy, dz);
            dx = fmax(0.0, fabs(dx) - hx); dy = fmax(0.0, fabs(dy) - hy); dz = fmax(0.0, fabs(dz) - hz);
            hit = hit || (dx * dx + dy * dy + dz * dz <= R * R);
        }
        return hit;
    }
#pragma unroll
    for (int g = 0; g < 2; g++) {
        if (g == 0 ? !use_old : !use_new) continue;
        const double dx = fmax(0.0, fabs(min_image_fast(m.cen[g][0] - bx, dc.L[0], dc.invL[0])) - hx);
        const double dy = fmax(0.0, fabs(min_image_fast(m.cen[g][1] - by, dc.L[1], dc.invL[1])) - hy);
        const double dz = fmax(0.0, fabs(min_image_fast(m.cen[g][2] - bz, dc.L[2], dc.invL[2])) - hz);
        const double R = m.cen[g][3];
        hit = hit || (dx * dx + dy * dy + dz * dz <= R * R);
    }
    return hit;
}

// ----------------------------------------------------------------------------------------------
// Real-space phase of one trial.  Work items = framework tiles in reach (ordered compaction of the cull
// result, so the item -> warp assignment and hence the summation order is deterministic) followed by tiles
// of 32 guest atoms.  Item i goes to global warp i mod (n_parts * warps).  All threads of the block.
// OLD / NEW: geometries this instantiation can evaluate; uo / un: run-time masks on top (true where the instantiation fixes the
// set); fw / guests: sweep the framework tiles / this chain's guests; cc: the sphere {x, y, z, R^2} of the geometries swept.
template <int NT, bool OLD, bool NEW, bool SHIFT, bool TRI>
__device__ __forceinline__ void pair_phase_t(const DevConst& dc, const ChainState& cs, int c, const SweepCtx& s, int part, int n_parts,
                                    bool uo, bool un, bool fw, bool guests, const double* cc, double (&acc)[4], unsigned& cnt, int& item0) {
    const MolS& m = *s.mol;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int gw = part * (NT / 32) + warp, gstride = n_parts * (NT / 32);
    PairConst pc = {dc.L[0], dc.L[1], dc.L[2], dc.invL[0], dc.invL[1], dc.invL[2], dc.reach2, dc.alpha, dc.rc2, dc.gtab_smin, dc.gtab_smax, dc.gtab_base, dc.gtab_deg_start[0], dc.gtab_deg_start[1]};
    const bool cull = dc.reach2 < 1e300;
    CompactQ cq = {0.0, 0.0, 0.0, 0.0, 0, 0};   // compacted sweep (SHIFT): survivors carried between tiles
    // item0 (in / out): items dealt so far in this trial -- a second call goes on dealing where the first one stopped
    for (int base = 0; fw && base < dc.n_tiles; base += CULL_CHUNK) {
        const int n = min(CULL_CHUNK, dc.n_tiles - base);
        const int nwords = (n + 31) >> 5;
        for (int i0 = warp * 32; i0 < n; i0 += NT) {
            const int i = i0 + lane;
            const bool hit = i < n && (!cull || tile_in_reach<TRI>(dc, m, base + i, OLD && uo, NEW && un));
            const unsigned b = __ballot_sync(0xffffffffu, hit);
            if (lane == 0) s.mask[i0 >> 5] = b;
        }
        __syncthreads();
        if (warp == 0) {  // exclusive prefix of the per-word counts
            int run = 0;
            for (int w0 = 0; w0 < nwords; w0 += 32) {
                const int w = w0 + lane;
                const int cnt = w < nwords ? __popc(s.mask[w]) : 0;
                int inc = cnt;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
                if (w < nwords) s.prefix[w] = run + inc - cnt;
                run += __shfl_sync(0xffffffffu, inc, 31);
            }
            if (lane == 0) s.mol->nlist = run;
        }
        __syncthreads();
        for (int i0 = warp * 32; i0 < n; i0 += NT) {
            const unsigned b = s.mask[i0 >> 5];
            if ((b >> lane) & 1u) s.list[s.prefix[i0 >> 5] + __popc(b & ((1u << lane) - 1u))] = (uint16_t)(i0 + lane);
        }
        __syncthreads();
        const int nlist = m.nlist;
#ifdef MMC_RS_SPLIT   // diagnostic (with MMC_PHASE_TIMING): slot 3 = tile cull, slot 5 = framework tiles, slot 2 = the rest of the real-space phase
        MMC_TSTAMP(3);
#endif
#ifdef MMC_COMPACT
        if (SHIFT) {
            // Compacted sweep: only ~half of the lanes of a visited tile hold an atom inside the molecule's sphere (tile boxes
            // are coarser than the sphere; r2f ncu: 23.8 of 32 threads active per instruction).  Each warp therefore shifts and
            // sphere-tests a tile's atoms (cheap), carries the survivors in registers and runs the expensive loop over the moved
            // atoms (4 x 2 pair evaluations with 9-coefficient table lookups each) only on FULL batches of 32 survivors.
            // Survivors keep their tile order, so the summation order stays deterministic.
            for (int it = (gw + gstride - item0 % gstride) % gstride; it < nlist; it += gstride) {
                const int j = (base + (int)s.list[it]) * TILE + lane;
                double xj = 0.0, yj = 0.0, zj = 0.0, qj = 0.0; int tj = 0; bool in = false;
                if (j < dc.n_fw) {
                    xj = dc.sx[j]; yj = dc.sy[j]; zj = dc.sz[j]; qj = dc.sq[j]; tj = dc.stype[j];
                    in = pair_shift<TRI>(dc, pc, cc, xj, yj, zj);
                }
                compact_push<OLD, NEW, TRI>(dc, pc, s, cq, in, xj, yj, zj, qj, tj, false, uo, un, cc, acc, cnt);
            }
        } else
#endif
#ifndef MMC_NO_TILE_PREFETCH
        {   // the next tile's atom is loaded (L2 latency) before this tile's pairs are evaluated
            int it = (gw + gstride - item0 % gstride) % gstride;
            double xn = 0.0, yn = 0.0, zn = 0.0, qn = 0.0; int tn = 0; bool vn = false;
            if (it < nlist) {
                const int j = (base + (int)s.list[it]) * TILE + lane;
                vn = j < dc.n_fw;
                if (vn) { xn = dc.sx[j]; yn = dc.sy[j]; zn = dc.sz[j]; qn = dc.sq[j]; tn = dc.stype[j]; }
            }
            while (it < nlist) {
                const double xj = xn, yj = yn, zj = zn, qj = qn; const int tj = tn; const bool v = vn;
                it += gstride;
                if (it < nlist) {
                    const int j = (base + (int)s.list[it]) * TILE + lane;
                    vn = j < dc.n_fw;
                    if (vn) { xn = dc.sx[j]; yn = dc.sy[j]; zn = dc.sz[j]; qn = dc.sq[j]; tn = dc.stype[j]; }
                }
                if (v) pair_atom<OLD, NEW, SHIFT, TRI>(dc, pc, s, dc.ntypes, xj, yj, zj, qj, tj, uo, un, cc, acc, cnt);
            }
        }
#else
        for (int it = (gw + gstride - item0 % gstride) % gstride; it < nlist; it += gstride) {
            const int j = (base + (int)s.list[it]) * TILE + lane;
            if (j < dc.n_fw) pair_atom<OLD, NEW, SHIFT, TRI>(dc, pc, s, dc.ntypes, dc.sx[j], dc.sy[j], dc.sz[j], dc.sq[j], dc.stype[j], uo, un, cc, acc, cnt);
        }
#endif
        item0 += nlist;
        __syncthreads();  // list is rebuilt by the next chunk
#ifdef MMC_RS_SPLIT
        MMC_TSTAMP(5);
#endif
    }
    // guests of this chain: slots [0, count) of every active residue, the moved slot excluded
    for (int ia = 0; guests && ia < dc.n_active; ia++) {
        const int r = dc.active_list[ia], na = dc.natoms[r];
        const int n_at = cs.count[c * MMC_MAX_RES + r] * na;
        const int ntile = (n_at + 31) >> 5;
        const size_t base = (size_t)c * dc.g_atoms + dc.g_atom_base[r];
        for (int it = (gw + gstride - item0 % gstride) % gstride; it < ntile; it += gstride) {
            const int i = it * 32 + lane;
#ifdef MMC_COMPACT
            if (SHIFT) {
                double xj = 0.0, yj = 0.0, zj = 0.0, qj = 0.0; int tj = 0; bool in = false;
                if (i < n_at) {
                    const int mol = i / na, a = i - mol * na;
                    if (!(m.excl && r == m.res && mol == m.mol)) {
                        xj = cs.gx[base + i]; yj = cs.gy[base + i]; zj = cs.gz[base + i];
                        qj = dc.res_q[r * MMC_MAX_MOL_ATOMS + a]; tj = dc.res_type[r * MMC_MAX_MOL_ATOMS + a];
                        in = pair_shift<TRI>(dc, pc, cc, xj, yj, zj);
                    }
                }
                compact_push<OLD, NEW, TRI>(dc, pc, s, cq, in, xj, yj, zj, qj, tj, false, uo, un, cc, acc, cnt);
                continue;
            }
#endif
            if (i < n_at) {
                const int mol = i / na, a = i - mol * na;
                if (!(m.excl && r == m.res && mol == m.mol))
                    pair_atom<OLD, NEW, SHIFT, TRI>(dc, pc, s, dc.ntypes, cs.gx[base + i], cs.gy[base + i], cs.gz[base + i], dc.res_q[r * MMC_MAX_MOL_ATOMS + a],
                                        dc.res_type[r * MMC_MAX_MOL_ATOMS + a], uo, un, cc, acc, cnt);
            }
        }
        item0 += ntile;
    }
#ifdef MMC_COMPACT
    if (SHIFT) compact_push<OLD, NEW, TRI>(dc, pc, s, cq, false, 0.0, 0.0, 0.0, 0.0, 0, true, uo, un, cc, acc, cnt);   // the partial last batch
#endif
}

// uo / un: geometries wanted; fw / guests: which atoms to sweep; cc: their sphere.  rt: use the one instantiation that masks
// OLD / NEW at run time (tilted cells always; the guests-only pass, which is small, so that it adds one inlined copy, not three)
template <int NT, bool TRI>
__device__ __forceinline__ void pair_phase(const DevConst& dc, const ChainState& cs, int c, const SweepCtx& s, int part, int n_parts,
                                  bool uo, bool un, bool fw, bool guests, bool rt, const double* cc, double (&acc)[4], unsigned& cnt, int& item0) {
    const MolS& m = *s.mol;
    if (m.kind <= 0 || !(uo || un) || !(fw || guests)) return;
    const bool sh = (cc == m.ccf ? m.shift_okf : m.shift_ok) && dc.reach2 < 1e300;
    if (TRI || rt) {
        if (sh) pair_phase_t<NT, true, true, true, TRI>(dc, cs, c, s, part, n_parts, uo, un, fw, guests, cc, acc, cnt, item0);
        else pair_phase_t<NT, true, true, false, TRI>(dc, cs, c, s, part, n_parts, uo, un, fw, guests, cc, acc, cnt, item0);
        return;
    }
    if (uo && un) { if (sh) pair_phase_t<NT, true, true, true, false>(dc, cs, c, s, part, n_parts, true, true, fw, guests, cc, acc, cnt, item0); else pair_phase_t<NT, true, true, false, false>(dc, cs, c, s, part, n_parts, true, true, fw, guests, cc, acc, cnt, item0); }
    else if (uo) { if (sh) pair_phase_t<NT, true, false, true, false>(dc, cs, c, s, part, n_parts, true, true, fw, guests, cc, acc, cnt, item0); else pair_phase_t<NT, true, false, false, false>(dc, cs, c, s, part, n_parts, true, true, fw, guests, cc, acc, cnt, item0); }
    else { if (sh) pair_phase_t<NT, false, true, true, false>(dc, cs, c, s, part, n_parts, true, true, fw, guests, cc, acc, cnt, item0); else pair_phase_t<NT, false, true, false, false>(dc, cs, c, s, part, n_parts, true, true, fw, guests, cc, acc, cnt, item0); }
}

// ----------------------------------------------------------------------------------------------
#ifdef MMC_RECIP_DMMA
// Reciprocal phase on the FP64 MMA unit.
//
// The k list is ordered (kx, ky, kz) with kz fastest, so all k of one (kx, ky) "row" are contiguous (kz = -m .. m; row (0, 0)
// lacks kz = 0) and the update of a row is a rank-n_term product
//     dS(row, kz) = sum_t P_t(row) Z_t(kz),   P_t = X_t(kx) Y_t(ky),   Z_t = q_t e^{i kz theta_z,t}.
// With the real and imaginary parts spread over the MMA's rows and columns, ONE pair of DMMA m8n8k4 instructions per 4 terms
// gives every lane the four real product sums of "its" (row, kz):
//     A operand (8 x 4) = Zr or Zi for 8 consecutive kz (rows) x 4 terms        -> fragment = 1 double per lane
//     B operand (4 x 8) = 4 terms x (4 rows x {Pr, Pi})                          -> fragment = 1 double per lane
//     C (8 x 8): lane (g = lane / 4, t = lane % 4) holds C[g][2t], C[g][2t + 1]  = sum_t Z(kz_g) . (Pr, Pi)(row_t)
//     with A = Zr: (s1, s4) = (sum Zr Pr, sum Zr Pi);   with A = Zi: (s3, s2) = (sum Zi Pr, sum Zi Pi)
// and sum P Z = (s1 - s2, s3 + s4) updates k(+kz), sum P conj(Z) = (s1 + s2, s4 - s3) updates k(-kz) (Z(-kz) = conj Z(kz)).
// The operands sit in register fragments: per 4 rows x 8 kz x (+-) = 64 k-vectors the warp issues four 8-byte shared loads
// (the A fragments) instead of two 16-byte loads per TERM per 2 k-vectors in the round-1 item loop, whose P / Z loads were
// 40 % of the kernel's shared-memory wavefronts (profiles/r1t_ncu_lines.txt).  The B fragments (the (kx, ky) products) are
// formed in registers from the 1-D tables once per 4-row group; nothing is staged per chunk and the phase has NO barrier:
// every warp walks its own row groups.  A is read and A' written exactly as before (A' goes to the other half of the
// double buffer); acc[0] += f W |A|^2, acc[1] += f W |A'|^2.
// The MMA is used as an operand-sharing device for a rank-4..8 update, not to reach tensor-core flops: the phase stays
// HBM-bound (40 B per k-eval) and tcgen05 has no FP64 kind.
__device__ __forceinline__ void dmma_884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// B fragment of K-block kb for the 4-row group starting at row rg: lane (g, t) supplies term 4 kb + t, row rg + g / 2,
// component g % 2 of P = X(kx) Y(ky)
__device__ __forceinline__ double recip_bfrag(const SweepCtx& s, int nterm, int term, bool row_ok, int ix, int iy, int comp) {
    if (!row_ok || term >= nterm) return 0.0;
    const double2 X = s.X[term * s.nx + ix], Y = s.Y[term * s.ny + iy];
    return comp ? fma(X.x, Y.y, X.y * Y.x) : fma(X.x, Y.x, -X.y * Y.y);
}

// KB > 0: the B fragments of KB K-blocks stay in registers for the whole group (n_term <= 4 KB); KB == 0: generic, any n_term,
// B fragments re-formed per kz block.  SPLIT: work unit = (group, kz block) instead of a whole group (few rows per CTA:
// strict mode and the k-sharded / cluster sweeps), otherwise a unit = one group and the kz blocks are walked inside.
template <int KB, bool TRI, bool SPLIT>
__device__ __forceinline__ void recip_groups(const DevConst& dc, const SweepCtx& s, const double2* __restrict__ A, double2* __restrict__ An,
                                             int r0, int r1, int unit0, int ustride, int nterm, double (&acc)[2]) {
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int ngroups = (r1 - r0 + 3) >> 2, nzb = s.nzb;
    const int nunits = SPLIT ? ngroups * nzb : ngroups;
    const int kbn = (nterm + 3) >> 2;
    const float inv_ny = 1.0f / (float)s.ny;
    // the k range of this lane's row is fetched one unit ahead (r2z: 3.9 % of the kernel's stall samples sat on these two
    // dependent loads at the head of every unit)
    int k0n = 0, k1n = 0;   // raw loads: nothing consumes them before the next unit starts (a difference formed here would wait for them)
    if (!SPLIT && unit0 < nunits && r0 + 4 * unit0 + t < r1) { k0n = dc.row_kstart[r0 + 4 * unit0 + t]; k1n = dc.row_kstart[r0 + 4 * unit0 + t + 1]; }
    for (int u = unit0; u < nunits; u += ustride) {
        const int q = SPLIT ? u / nzb : u;
        const int rg = r0 + 4 * q;
        // the row this lane updates: k range [k0, k1) = kz -m .. m
        const int r = rg + t;
        int k0 = 0, n = 0;
        if (SPLIT) {
            if (r < r1) { k0 = dc.row_kstart[r]; n = dc.row_kstart[r + 1] - k0; }
        } else {
            k0 = k0n; n = k1n - k0n;
            const int rr = r + 4 * ustride;
            k0n = 0; k1n = 0;
            if (u + ustride < nunits && rr < r1) { k0n = dc.row_kstart[rr]; k1n = dc.row_kstart[rr + 1]; }
        }
        const int m = n >> 1, has0 = n & 1;
        const int mmax = __reduce_max_sync(0xffffffffu, n > 0 ? m : -1);
        if (mmax < 0) continue;                                   // four empty rows (outside the k sphere)
        const int jlo = SPLIT ? u - q * nzb : 0, jhi = SPLIT ? jlo + 1 : (mmax >> 3) + 1;
        if (SPLIT && 8 * jlo > mmax) continue;
        const int base_p = k0 + m - (1 - has0);                   // k(+kz) = base_p + kz, k(-kz) = k0 + m - kz
        // Software pipeline over the kz blocks: the A(k) / f W loads of block j + 1 are issued BEFORE block j is computed, so
        // two blocks' worth of memory requests are in flight per warp.  (r2c phase timing: 1 090 cycles per block with one
        // block in flight -- the whole DRAM / L2 latency exposed 170 times per trial and warp.)
        const double2 zero2 = make_double2(0.0, 0.0);
        int kz = 8 * jlo + g;
        bool ok = kz <= m && (kz >= 1 || has0);
        double2 ap = ok ? A[base_p + kz] : zero2;
        double2 am = (ok && kz != 0) ? A[k0 + m - kz] : zero2;
        double w = ok ? dc.fw[base_p + kz] : 0.0;   // f W depends on kx and |k|^2 only: the same for +kz and -kz in an orthorhombic cell
        double wm = (TRI && ok && kz != 0) ? dc.fw[k0 + m - kz] : w;   // a tilted cell has |k(+kz)| != |k(-kz)|
        // B fragments: row rg + g / 2, component g % 2
        const int rb = rg + (g >> 1);
        const bool rb_ok = rb < r1;
        // rb / ny without the integer-division sequence: (rb + 1/2) / ny is at least 1 / (2 ny) away from an integer, the
        // single-precision error of a quotient below nx (a few hundred at most) is orders of magnitude smaller
        const int ix = rb_ok ? __float2int_rd(((float)rb + 0.5f) * inv_ny) : 0, iy = rb_ok ? rb - ix * s.ny : 0;
        double bf[KB > 0 ? KB : 1];
        if (KB > 0) {
#pragma unroll
            for (int kb = 0; kb < KB; kb++) bf[kb] = recip_bfrag(s, nterm, 4 * kb + t, rb_ok, ix, iy, g & 1);
        }
        if (!SPLIT) {
            // pull the group this warp takes next from HBM into L2 while this one is computed: one bulk prefetch (no LSU
            // data wavefronts, unlike a prefetch per line).  Its k range [ka, kb) comes from the row ranges lanes 0..3 just
            // fetched for the next unit (rows are contiguous in k, rows beyond r1 hold (0, 0))
            int kb = k1n;
            kb = max(kb, __shfl_xor_sync(0xffffffffu, kb, 1)); kb = max(kb, __shfl_xor_sync(0xffffffffu, kb, 2));
            if (lane == 0 && kb > k0n) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(A + k0n), "r"((kb - k0n) * 16) : "memory");
        }
        for (int j = jlo; j < jhi; j++) {
            const int kzn = kz + 8;                                  // >= 8: never the kz = 0 column
            const bool okn = j + 1 < jhi && kzn <= m;
            double2 apn = zero2, amn = zero2; double wn = 0.0, wmn = 0.0;
            if (okn) {
                apn = A[base_p + kzn]; amn = A[k0 + m - kzn]; wn = dc.fw[base_p + kzn];
                wmn = TRI ? dc.fw[k0 + m - kzn] : wn;
            }
            double s1 = 0.0, s4 = 0.0, s3 = 0.0, s2 = 0.0;
            const double* zr = s.ZR + t * s.zS + 8 * j + g;
            const double* zi = s.ZI + t * s.zS + 8 * j + g;
            if (KB > 0) {
#pragma unroll
                for (int kb = 0; kb < KB; kb++) {
                    dmma_884(s1, s4, zr[4 * kb * s.zS], bf[kb]);
                    dmma_884(s3, s2, zi[4 * kb * s.zS], bf[kb]);
                }
            } else {
                for (int kb = 0; kb < kbn; kb++) {
                    const double b = recip_bfrag(s, nterm, 4 * kb + t, rb_ok, ix, iy, g & 1);
                    dmma_884(s1, s4, zr[4 * kb * s.zS], b);
                    dmma_884(s3, s2, zi[4 * kb * s.zS], b);
                }
            }
            if (ok) {
                const int kpl = base_p + kz, kmi = k0 + m - kz;
                const double2 bp = make_double2(ap.x + (s1 - s2), ap.y + (s3 + s4));
                An[kpl] = bp;
                acc[0] = fma(w, fma(ap.x, ap.x, ap.y * ap.y), acc[0]);  // amplitude_squared helper_utils.f90:125
                acc[1] = fma(w, fma(bp.x, bp.x, bp.y * bp.y), acc[1]);
                if (kz != 0) {
                    const double2 bm = make_double2(am.x + (s1 + s2), am.y + (s4 - s3));
                    An[kmi] = bm;
                    acc[0] = fma(wm, fma(am.x, am.x, am.y * am.y), acc[0]);
                    acc[1] = fma(wm, fma(bm.x, bm.x, bm.y * bm.y), acc[1]);
                }
            }
            ap = apn; am = amn; w = wn; wm = wmn; kz = kzn; ok = okn;
        }
    }
}

template <int NT, bool TRI>
__device__ __forceinline__ void recip_phase(const DevConst& dc, const ChainState& cs, int c, const SweepCtx& s, int part, int n_parts,
                                   double (&acc)[2]) {
    const MolS& m = *s.mol;
    if (m.kind <= 0) return;
    const int nterm = m.nterm;
    if (nterm == 0) {
        // no charged atom moved: A' = A.  The energies are still summed and A copied (the reference re-sums it, ewald_energy.f90:259-272)
    }
    const int cur = __shfl_sync(0xffffffffu, m.cur, 0);
    const double2* A = cs.A + ((size_t)c * 2 + cur) * dc.nk;
    double2* An = cs.A + ((size_t)c * 2 + (cur ^ 1)) * dc.nk;
    const int nrows = s.nx * s.ny;
    const int rpp = (nrows + n_parts - 1) / n_parts;
    const int r0 = min(nrows, part * rpp), r1 = min(nrows, r0 + rpp);
    if (r0 >= r1) return;
    const int warp = threadIdx.x >> 5;
    if (n_parts > 1 && (r1 - r0) < 8 * (NT / 32)) {   // fewer than two 4-row groups per warp (strict mode, k-sharded sweeps): spread (group, kz block) units over the warps
        if (nterm <= 4) recip_groups<1, TRI, true>(dc, s, A, An, r0, r1, warp, NT / 32, nterm, acc);
        else if (nterm <= 8) recip_groups<2, TRI, true>(dc, s, A, An, r0, r1, warp, NT / 32, nterm, acc);
        else recip_groups<0, TRI, true>(dc, s, A, An, r0, r1, warp, NT / 32, nterm, acc);
    } else {
        if (nterm <= 4) recip_groups<1, TRI, false>(dc, s, A, An, r0, r1, warp, NT / 32, nterm, acc);
        else if (nterm <= 8) recip_groups<2, TRI, false>(dc, s, A, An, r0, r1, warp, NT / 32, nterm, acc);
        else recip_groups<0, TRI, false>(dc, s, A, An, r0, r1, warp, NT / 32, nterm, acc);
    }
}
#else   // MMC_RECIP_ITEMS: the round-1 item loop
// Reciprocal phase.  The k list is ordered (kx, ky, kz) with kz fastest, so all k of one (kx, ky) "row" are
// contiguous and share the product X_t(kx) Y_t(ky).  The rows (flattened index kx*ny + ky+kmax_y) are walked in
// chunks of NT/4: the chunk's products are staged in shared memory (double buffered: the next chunk is staged
// while this one is consumed, one barrier per chunk).  One work item per thread per pass = the two k-vectors
// (kx, ky, +kz) and (kx, ky, -kz): Z(-kz) = conj Z(kz), so both updates come from the same four real product sums
// (4 FMAs and two 16-byte shared loads per term for TWO k-vectors).  A is read ascending for +kz and descending for
// -kz (both coalesced), A' goes to the other half of the double buffer.  acc[0] += f W |A|^2, acc[1] += f W |A'|^2.
template <int NT>
__device__ __forceinline__ void recip_stage_rows(const SweepCtx& s, int nterm, int r_a, int r_b, double2* buf) {
    const int nrow = r_b - r_a;
    for (int i = threadIdx.x; i < nrow * nterm; i += NT) {
        const int t = i / nrow, rl = i - t * nrow;
        const int r = r_a + rl, ix = r / s.ny, iy = r - ix * s.ny;
        const double2 X = s.X[t * s.nx + ix], Y = s.Y[t * s.ny + iy];
        buf[rl * s.zs + t] = make_double2(fma(X.x, Y.x, -X.y * Y.y), fma(X.x, Y.y, X.y * Y.x));
    }
}

// the items [i0, i1) of one staged chunk of rows.  NTERM > 0: term count known at compile time (the shared loads get
// immediate offsets and the loop disappears); NTERM == 0: generic loop over nterm.
template <int NT, int NTERM, bool TRI, bool FACT = false>
__device__ __forceinline__ void recip_items(const DevConst& dc, const SweepCtx& s, const double2* __restrict__ A, double2* __restrict__ An,
                                            const double2* xy, int ra, int i0, int i1, int nterm, double (&acc)[2]) {
    const int ny = s.ny, zs = s.zs;
    int2 it_next = (i0 + (int)threadIdx.x < i1) ? dc.kitem[i0 + threadIdx.x] : make_int2(0, 0);
    for (int i = i0 + threadIdx.x; i < i1; i += NT) {
        const int2 it = it_next;
        if (i + NT < i1) it_next = dc.kitem[i + NT];
        const uint32_t kp = (uint32_t)it.x;
        const int ix = kp & 1023u, iy = (kp >> 10) & 2047u, kz = (kp >> 21) & 1023u;
        const int kpl = it.y, kmi = kpl - 2 * kz + (int)(kp >> 31);
        const bool two = kz != 0;
        const double2 ap = A[kpl];
        const double2 am = two ? A[kmi] : make_double2(0.0, 0.0);
        const double w = dc.fw[kpl];   // f W depends on kx and |k|^2 only: the same for +kz and -kz in an orthorhombic cell
        const double2* pt = xy + (ix * ny + iy - ra) * zs;
        const double2* zt = s.Z + kz * zs;
        // sum_t P_t Z_t and sum_t P_t conj(Z_t) from the four real product sums
        double s1 = 0.0, s2 = 0.0, s3 = 0.0, s4 = 0.0;
        if (NTERM > 0) {
#pragma unroll
            for (int t = 0; t < (FACT ? NTERM - 1 : NTERM); t++) {
                const double2 P = pt[t], Z = zt[t];
                s1 = fma(P.x, Z.x, s1); s2 = fma(P.y, Z.y, s2);
                s3 = fma(P.x, Z.y, s3); s4 = fma(P.y, Z.x, s4);
            }
        } else {
            for (int t = 0; t < (FACT ? nterm - 1 : nterm); t++) {
                const double2 P = pt[t], Z = zt[t];
                s1 = fma(P.x, Z.x, s1); s2 = fma(P.y, Z.y, s2);
                s3 = fma(P.x, Z.y, s3); s4 = fma(P.y, Z.x, s4);
            }
        }
        double dpx = s1 - s2, dpy = s3 + s4, dmx = s1 + s2, dmy = s4 - s3;   // sum_t P_t Z_t and sum_t P_t conj(Z_t)
        if (FACT) {   // times (e^{i k d} - 1), the last term's product, for k(+kz) and k(-kz)
            const double2 P = pt[(NTERM > 0 ? NTERM : nterm) - 1], Z = zt[(NTERM > 0 ? NTERM : nterm) - 1];
            const double d1 = P.x * Z.x, d2 = P.y * Z.y, d3 = P.x * Z.y, d4 = P.y * Z.x;
            const double epx = (d1 - d2) - 1.0, epy = d3 + d4, emx = (d1 + d2) - 1.0, emy = d4 - d3;
            const double tpx = fma(dpx, epx, -dpy * epy), tmx = fma(dmx, emx, -dmy * emy);
            dpy = fma(dpx, epy, dpy * epx); dmy = fma(dmx, emy, dmy * emx);
            dpx = tpx; dmx = tmx;
        }
        const double2 bp = make_double2(ap.x + dpx, ap.y + dpy);
        An[kpl] = bp;
        acc[0] = fma(w, fma(ap.x, ap.x, ap.y * ap.y), acc[0]);  // amplitude_squared helper_utils.f90:125
        acc[1] = fma(w, fma(bp.x, bp.x, bp.y * bp.y), acc[1]);
        if (two) {
            const double2 bm = make_double2(am.x + dmx, am.y + dmy);
            An[kmi] = bm;
            const double wm = TRI ? dc.fw[kmi] : w;   // a tilted cell has |k(+kz)| != |k(-kz)|
            acc[0] = fma(wm, fma(am.x, am.x, am.y * am.y), acc[0]);
            acc[1] = fma(wm, fma(bm.x, bm.x, bm.y * bm.y), acc[1]);
        }
    }
}

template <int NT, bool TRI>
__device__ __forceinline__ void recip_phase(const DevConst& dc, const ChainState& cs, int c, const SweepCtx& s, int part, int n_parts,
                                   double (&acc)[2]) {
    constexpr int RMAX = NT / 4;
    const MolS& m = *s.mol;
    if (m.kind <= 0) return;
    const int nterm = m.nterm;
    const double2* A = cs.A + ((size_t)c * 2 + m.cur) * dc.nk;
    double2* An = cs.A + ((size_t)c * 2 + (m.cur ^ 1)) * dc.nk;
    const int nrows = s.nx * s.ny;
    const int rpp = (nrows + n_parts - 1) / n_parts;
    const int r0 = min(nrows, part * rpp), r1 = min(nrows, r0 + rpp);
    if (r0 >= r1) return;
    const size_t bufsz = (size_t)RMAX * s.zs;
    int buf = 0;
    recip_stage_rows<NT>(s, nterm, r0, min(r1, r0 + RMAX), s.XY);
    __syncthreads();
    for (int ra = r0; ra < r1; ra += RMAX) {
        const int rb = min(r1, ra + RMAX);
        if (rb < r1) {
            recip_stage_rows<NT>(s, nterm, rb, min(r1, rb + RMAX), s.XY + (buf ^ 1) * bufsz);
            // pull the next chunk's A(k) from HBM into L2 while this chunk is computed: ONE bulk prefetch by one thread
            // (no LSU wavefronts, unlike a prefetch per line)
            if (threadIdx.x == 0) {
                const int ka = dc.row_kstart[rb], kb = dc.row_kstart[min(r1, rb + RMAX)];
                if (kb > ka) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(A + ka), "r"((kb - ka) * 16) : "memory");
            }
        }
        const double2* xy = s.XY + buf * bufsz;
        const int i0 = dc.item_row_start[ra], i1 = dc.item_row_start[rb];
#ifdef MMC_FACTORED_TRANSLATION
        if (m.fact) {
            switch (nterm) {   // charged atoms + 1
                case 3: recip_items<NT, 3, TRI, true>(dc, s, A, An, xy, ra, i0, i1, nterm, acc); break;
                case 4: recip_items<NT, 4, TRI, true>(dc, s, A, An, xy, ra, i0, i1, nterm, acc); break;
                case 5: recip_items<NT, 5, TRI, true>(dc, s, A, An, xy, ra, i0, i1, nterm, acc); break;
                default: recip_items<NT, 0, TRI, true>(dc, s, A, An, xy, ra, i0, i1, nterm, acc); break;
            }
        } else
#endif
        switch (nterm) {   // 3 / 6 = rigid water insertion-deletion / move; the others cover 1..4 charged sites
            case 1: recip_items<NT, 1, TRI>(dc, s, A, An, xy, ra, i0, i1, nterm, acc); break;
            case 2: recip_items<NT, 2, TRI>(dc, s, A, An, xy, ra, i0, i1, nterm, acc); break;
            case 3: recip_items<NT, 3, TRI>(dc, s, A, An, xy, ra, i0, i1, nterm, acc); break;
            case 4: recip_items<NT, 4, TRI>(dc, s, A, An, xy, ra, i0, i1, nterm, acc); break;
            case 6: recip_items<NT, 6, TRI>(dc, s, A, An, xy, ra, i0, i1, nterm, acc); break;
            case 8: recip_items<NT, 8, TRI>(dc, s, A, An, xy, ra, i0, i1, nterm, acc); break;
            default: recip_items<NT, 0, TRI>(dc, s, A, An, xy, ra, i0, i1, nterm, acc); break;
        }
        __syncthreads();
        buf ^= 1;
    }
}

#endif  // MMC_RECIP_DMMA

// After sweep_trial(allow_cache) and the sum over the CTAs that shared the trial: v[0..3] = the guests' part of the real-space
// sums, fw[0..3] = the framework's.  Puts them together (the old geometry's framework part from the cache when it was known)
// and leaves the cache's next values in the Trial for commit_trial.  Single thread.
__device__ __forceinline__ void fold_fw_cache(Trial& t, const MolS& m, double* v, const double* fw) {
    if (!m.cache_on) return;
    const bool cached = m.fw_cached != 0;
    const double fo0 = cached ? m.fw_old[0] : fw[0], fo1 = cached ? m.fw_old[1] : fw[1];
    v[0] += fo0; v[1] += fo1; v[2] += fw[2]; v[3] += fw[3];
    t.fw_new_ok = m.do_new != 0; t.fw_new[0] = fw[2]; t.fw_new[1] = fw[3];
    t.fw_old_fresh = m.fo != 0; t.fw_old[0] = fo0; t.fw_old[1] = fo1;
}

// One trial of chain c, all threads of the CTA: stage, real-space sweep(s), reciprocal sweep.  `part` of `n_parts`
// CTAs share the trial (1 of 1 in batched mode).  phases: bit 0 real space, bit 1 reciprocal.
// Result: s.mol->sums = {lj_old, coul_old, lj_new, coul_new, sum fW|A|^2, sum fW|A'|^2} of this CTA (block sums in a
// fixed order), valid after the trailing barrier.  The real-space sums are reduced and parked in shared memory BEFORE
// the reciprocal sweep so that its item loop does not carry them in registers (64 registers per thread at 4 CTAs/SM).
// TRI: compiled for tilted cells (dc.tri); the orthorhombic kernels carry none of that code
template <int NT, bool TRI>
__device__ __forceinline__ void sweep_trial(const DevConst& dc, const ChainState& cs, int c, const SweepCtx& s, const Trial& t, int part, int n_parts, int phases,
                                            bool allow_cache = false) {
    // allow_cache (the batched kernels, which own the Trial and commit it themselves): the framework-energy cache is read; the
    // caller then runs fold_fw_cache on the (cluster-wide) sums
    const bool swap = t.kind == MMC_SWAP;
    const bool cache = allow_cache && dc.fw_cache && !swap && (phases & 1);
    stage_terms<TRI>(dc, cs, c, s, t);
    stage_molecule(dc, cs, c, s, t, swap ? 1 : 0, cache);
    MMC_TSTAMP(1);
    {
        // pa[0..3]: {LJ, Coulomb} of the old and of the new geometry.  With the cache the framework part is summed on its own, so
        // that the new geometry's can be stored (the old geometry's is skipped when it is known), and parked as rows 4..7 of
        // s.red before the guests are swept: the two sets of partial sums never hold registers at the same time
        double pa[4] = {0, 0, 0, 0};
        unsigned cnt = 0u;   // executed work of this thread in this trial: low 16 bits = distances formed, high 16 = pairs inside reach
        int item0 = 0;       // work items dealt to the warps so far (the guests' tiles go on where the framework's stopped)
        const MolS& m = *s.mol;
        // one call site per sweep flavour (each is a fully inlined copy of the pair loop): a swap runs it twice, first for the
        // removed molecule, then for the inserted one (different atom set, the removed one still excluded)
        for (int pass = 0; pass < (swap ? 2 : 1); pass++) {
            if (pass) { __syncthreads(); stage_molecule(dc, cs, c, s, t, 2); }
            if (!(phases & 1)) continue;
            // framework (and, without the cache, the guests as well): m.fo / m.ccf equal do_old / cc when nothing is cached
            pair_phase<NT, TRI>(dc, cs, c, s, part, n_parts, m.fo != 0, m.do_new != 0, true, !cache, false, m.ccf, pa, cnt, item0);
            if (cache) {
                block_sum_park<4, NT>(pa, s.red, 4);
                for (int i = 0; i < 4; i++) pa[i] = 0.0;
                pair_phase<NT, TRI>(dc, cs, c, s, part, n_parts, m.do_old != 0, m.do_new != 0, false, true, true, m.cc, pa, cnt, item0);
            }
        }
        const unsigned w0 = __reduce_add_sync(0xffffffffu, cnt & 0xffffu), w1 = __reduce_add_sync(0xffffffffu, cnt >> 16);
        if ((threadIdx.x & 31) == 0 && (w0 | w1)) { atomicAdd(&s.mol->work[0], w0); atomicAdd(&s.mol->work[1], w1); }
#ifdef MMC_PHASE_TIMING
        __syncthreads(); MMC_TSTAMP(2);
#endif
        if (cache) {
            block_sum_park<4, NT>(pa, s.red, 0);
            double v[8];
            block_sum_gather<8, NT>(v, s.red);
            if (threadIdx.x == 0) {
                for (int i = 0; i < 4; i++) { s.mol->sums[i] = v[i]; s.mol->fws[i] = v[4 + i]; }
                s.mol->cache_on = 1;
            }
        } else {
            block_sum<4, NT>(pa, s.red);
            if (threadIdx.x == 0) { for (int i = 0; i < 4; i++) s.mol->sums[i] = pa[i]; s.mol->cache_on = 0; }
        }
        MMC_TSTAMP(3);
    }
    {
        double ra[2] = {0, 0};
        if (phases & 2) recip_phase<NT, TRI>(dc, cs, c, s, part, n_parts, ra);
#ifdef MMC_PHASE_TIMING
        __syncthreads(); MMC_TSTAMP(4);
#endif
        block_sum<2, NT>(ra, s.red);
        if (threadIdx.x == 0) { s.mol->sums[4] = ra[0]; s.mol->sums[5] = ra[1]; }
    }
    __syncthreads();
    MMC_TSTAMP(5);
}

// executed-work counters of the real-space sweep (bench.py's executed-flop figure): cs.pair_count[0] += distances
// formed, [1] += pairs inside reach (Coulomb table and/or LJ evaluated).  One atomic pair per CTA per launch.
__device__ __forceinline__ void pair_count_init(const SweepCtx& s) {
    if (threadIdx.x == 0) {
        s.mol->work[0] = 0u; s.mol->work[1] = 0u;
#ifdef MMC_PHASE_TIMING
        for (int i = 0; i < 8; i++) s.mol->phase[i] = 0ull;
        s.mol->t_last = clock64();
#endif
    }
}
__device__ __forceinline__ void pair_count_flush(const ChainState& cs, const SweepCtx& s) {
    __syncthreads();
    if (threadIdx.x == 0 && (s.mol->work[0] | s.mol->work[1])) {
        atomicAdd(cs.pair_count, (unsigned long long)s.mol->work[0]);
        atomicAdd(cs.pair_count + 1, (unsigned long long)s.mol->work[1]);
    }
#ifdef MMC_PHASE_TIMING
    if (threadIdx.x == 0) for (int i = 0; i < 8; i++) atomicAdd(cs.pair_count + 2 + i, s.mol->phase[i]);
#endif
}

// strict mode / sharded step: grid = (dc.sweep_parts, chains) CTAs; the sweep_parts CTAs of one row share the pending
// trial of chain c0 + blockIdx.y; partial sums -> cs.part.  With a k/tile-sharded handle (dc.shard_count > 1, one
// handle per GPU) this rank's CTAs are parts [shard_rank * sweep_parts, (shard_rank + 1) * sweep_parts) of
// shard_count * sweep_parts: it visits 1/shard_count of the tiles and owns 1/shard_count of the (kx, ky) rows of A(k).
__global__ void __launch_bounds__(SWEEP_THREADS) k_sweep(DevConst dc, ChainState cs, int c0) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int c = c0 + blockIdx.y;
    const int part = dc.shard_rank * gridDim.x + blockIdx.x, n_parts = dc.shard_count * gridDim.x;
    const SweepCtx s = sweep_carve(dc, smem, SWEEP_THREADS);
    sweep_load_tables(dc, s);
    pair_count_init(s);
    if (dc.tri) sweep_trial<SWEEP_THREADS, true>(dc, cs, c, s, cs.trial[c], part, n_parts, 3);
    else sweep_trial<SWEEP_THREADS, false>(dc, cs, c, s, cs.trial[c], part, n_parts, 3);
    if (threadIdx.x == 0) {
        double* out = cs.part + ((size_t)c * dc.sweep_parts + blockIdx.x) * 6;
        for (int i = 0; i < 6; i++) out[i] = s.mol->sums[i];
    }
    pair_count_flush(cs, s);
}

// ---- strict mode, fused: ONE launch per trial ------------------------------------------------------------------------
// grid = dc.sweep_parts CTAs sharing the trial of one chain.  Every CTA builds the Trial in its own shared memory from the
// host's proposal (kernel argument: no H2D copy, no separate set-up launch), sweeps its share of the tiles and of the (kx, ky)
// rows and leaves six partial sums in cs.part; the LAST CTA to finish (ticket counter) adds the partials in a fixed order,
// assembles old / new energies and the acceptance probability (ComputeOldEnergy / ComputeNewEnergy /
// mc_acceptance_probability, monte_carlo_utils.f90:184-396) and writes them straight into host-mapped pinned memory, followed
// by a sequence number the host spins on: no device-to-host copy, no stream synchronisation.  decide != 0: the host passed the
// accept uniform it drew for this move (the Fortran move routines draw it right after the energies, so drawing it before the
// call leaves the stream's order unchanged) and the same CTA commits (AcceptMove / Reject*, :410-510) -- one launch per move.
// decide == 0: the Trial is parked in cs.trial[c] for a later mmc_trial_commit (k_commit).
struct StrictOut { double old_e[MMC_NE], new_e[MMC_NE], prob; int accepted, error; unsigned long long seq; };

__global__ void __launch_bounds__(SWEEP_THREADS) k_strict_trial(DevConst dc, ChainState cs, HostTrial ht, int decide, double u,
                                                                unsigned int* ticket, StrictOut* out, unsigned long long seq) {
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ int s_last;
    const int c = ht.chain, tid = threadIdx.x;
    const int part = blockIdx.x, n_parts = gridDim.x;
    const SweepCtx s = sweep_carve(dc, smem, SWEEP_THREADS);
    Trial& t = *s.trial;
    sweep_load_tables(dc, s);
    pair_count_init(s);
    if (tid == 0) {
        t.error = 0; t.accepted = 0; t.draw = 0; t.prob = 0.0;
        for (int i = 0; i < MMC_NE; i++) { t.old_e[i] = 0.0; t.new_e[i] = 0.0; }
        build_trial(dc, cs, c, t, ht.kind, ht.res, ht.mol, ht.com, ht.off, ht.res2, blockIdx.x == 0);
    }
    __syncthreads();
    if (dc.tri) sweep_trial<SWEEP_THREADS, true>(dc, cs, c, s, t, part, n_parts, 3);
    else sweep_trial<SWEEP_THREADS, false>(dc, cs, c, s, t, part, n_parts, 3);
    if (tid == 0) {
        double* o = cs.part + ((size_t)c * dc.sweep_parts + blockIdx.x) * 6;
        for (int i = 0; i < 6; i++) o[i] = s.mol->sums[i];
        __threadfence();
        s_last = atomicAdd(ticket, 1u) == (unsigned)(gridDim.x - 1);
    }
    pair_count_flush(cs, s);
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double v[6] = {0, 0, 0, 0, 0, 0};
    const double* pp = cs.part + (size_t)c * dc.sweep_parts * 6;
    for (int i = tid; i < n_parts; i += SWEEP_THREADS)
        for (int k = 0; k < 6; k++) v[k] += __ldcg(pp + i * 6 + k);
    block_sum<6, SWEEP_THREADS>(v, s.red);
    if (tid == 0) {
        finalize_trial(dc, cs, c, t, v);
        int acc = -1;
        if (decide) { acc = u <= t.prob ? 1 : 0; commit_trial(dc, cs, c, t, acc); }
        else cs.trial[c] = t;
        for (int i = 0; i < MMC_NE; i++) { out->old_e[i] = t.old_e[i]; out->new_e[i] = t.new_e[i]; }
        out->prob = t.prob; out->accepted = acc; out->error = 0;
        *ticket = 0u;
        __threadfence_system();
        *(volatile unsigned long long*)&out->seq = seq;
    }
}
