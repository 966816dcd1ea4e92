// common.cuh -- device-side data layout and helpers of libmaniacgpu (sm_100a).
//
// HBM layout (see DESIGN.md "Data layout"):
//   shared, read-only : framework SoA fx/fy/fz/fq (double) + ftype/fmol (int) for the single-point kernels and, for the
//                       per-move sweeps, the same atoms in Morton order in tiles of 32 (sx/sy/sz/sq/stype, tbox);
//                       LJ table lj[ti*ntypes+tj] = {sigma, 4*epsilon}; k-vector list kvec (int4) and the reciprocal work
//                       items kitem (one per (kx, ky) row and kz >= 0); fw[k] = form_factor(k) * W(k)
//   per chain, mutable: guest SoA gx/gy/gz (absolute positions, slot-major), gcom/goff (COM + site offsets, the
//                       reference's mol_com/site_offset), count, A(k) double-buffered (double2, 2*K per chain) with a
//                       parity bit, tracked energies, counters, one Trial descriptor, a Philox stream id.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/maniac_gpu.h"

#define MMC_MAX_RES 8
#define RECIP_THREADS 256   // single-point S(k) kernels
#define TAIL_THREADS 128
#define SWEEP_THREADS 256   // strict-mode sweep: CTAs of this size, dc.sweep_parts per chain
#define CLUSTER_THREADS 512 // k_chain_run_cl: CTA size of the cluster-per-chain kernel (few chains)
#define TILE 32             // framework atoms per tile (one per lane)
#define CULL_CHUNK 4096     // tiles culled per pass (shared-memory list)

// constants.f90:7-20
#define MMC_PI 3.14159265358979323846
#define MMC_TWOPI (2.0 * MMC_PI)
#define MMC_EPS0_INV_eVA 14.40198
#define MMC_KB_eVK 8.6173852e-5
#define MMC_ERR 1.0e-10
#define MMC_PROB_CREATE_DELETE 0.5  // parameters.f90:21
// bits of ChainState.error_flag: sticky on the device until a host entry point reads (and clears) them
#define MMC_ERRF_CAPACITY 1   // a creation / swap hit the slot capacity (CheckMoleculeIndex, output_utils.f90:604); the move was skipped
#define MMC_ERRF_TAPE 2       // a chain consumed more uniforms than its tape holds

struct DevConst {
    int nres, ntypes, n_chains, n_active;
    int natoms[MMC_MAX_RES], is_active[MMC_MAX_RES], cap[MMC_MAX_RES], active_list[MMC_MAX_RES];
    int g_atom_base[MMC_MAX_RES];  // first guest-atom index of residue r inside a chain's guest arrays
    int g_mol_base[MMC_MAX_RES];   // first guest-molecule slot of residue r
    int g_atoms, g_mols;           // per-chain guest capacities (atoms, molecules)
    int n_fw;                      // framework atoms (all inactive residues, flattened)
    int n_k_tiles;                 // ceil(nk / RECIP_THREADS)
    int nk, kmax[3], kstride;      // kstride = max(kmax)+1 : row length of a 1-D phase table
    int deletion_mode, use_table;
    int fw_cache;                  // 1: cache the framework part of every guest molecule's pair energy (ChainState.fw_pair)
    double q[MMC_MAX_RES][MMC_MAX_MOL_ATOMS];  // active residues only
    int type[MMC_MAX_RES][MMC_MAX_MOL_ATOMS];  // 0-based atom types
    double self_e[MMC_MAX_RES];                // ComputeEwaldSelfInteractionSingleMol per residue (K)
    double lo[3], L[3], invL[3], volume;
    // triclinic cell a = (L0,0,0), b = (xy,L1,0), c = (xz,yz,L2): tilt = {xy, xz, yz}; tri = any tilt non-zero;
    // half_min2 = (min(L)/2)^2: a brick image shorter than that is the nearest image
    double tilt[3], half_min2; int tri;
    double alpha, rc, coul_const, recip_const;  // coul_const = EPS0_INV_eVA/KB_eVK ; recip_const = coul_const*2pi/V
    double p_translation, p_rotation, p_swap;
    // pointers
    const double *fx, *fy, *fz, *fq;
    const int *ftype, *fmol;
    const double2* lj;
    const int4* kvec;
    const double* fw;
    const double *tab_erfc, *tab_r6, *tab_r12;  // tabulated_utils.f90 tables (5001 points each)
    double tab_dx;
    // ---- per-move sweep layout (kernels_sweep.cuh) ----
    // framework atoms in Morton order of their cell, padded to whole tiles of 32 atoms; tile t owns
    // atoms [32 t, 32 t + 32).  tbox[c][t], c = centre xyz, half extent xyz of the tile's wrapped positions (SoA: coalesced in the cull).
    const double *sx, *sy, *sz, *sq;
    const int* stype;
    const double* tbox;            // [6][n_tiles]
    int n_tiles;
    const double* gtab;            // [(MMC_GTAB_DEG+1)][MMC_GTAB_MAXSEG] Coulomb kernel table for this alpha
    int gtab_base;                 // segment index of s = gtab_smin
    int gtab_deg_start[2];         // segments >= [0] are fitted with degree D - 1, >= [1] with D - 2 (their top coefficients are not loaded)
    double gtab_smin, gtab_smax;   // table domain in r^2; outside it coul_kernel_slow is used
    // reciprocal work items (build_sweep_layout): one per (kx, ky) row and kz >= 0, serving k(+kz) and k(-kz).
    // .x = kx | (ky+kmax_y)<<10 | kz<<21 | flag<<31, .y = index of k(+kz); index of k(-kz) = .y - 2 kz + flag
    const int2* kitem;
    const int* item_row_start;     // [nx*ny + 1] first item of every flattened (kx, ky) row
    const int* row_kstart;         // [nx*ny + 1] first k-vector of every row (the k list is kz-fastest)
    int n_items;
    const double* res_q;           // [MMC_MAX_RES][MMC_MAX_MOL_ATOMS] charges of the active residues
    const int* res_type;           // same, 0-based atom types
    // real-space reach: LJ inside rc, erfc-Coulomb inside x_cut/alpha (terms beyond it are < erfc(x_cut),
    // far below one ulp of the sums); reach2 = inf when culling is off (every pair evaluated, Q3 literally)
    double rc2, alpha2, reach, reach2;
    int max_nat, sweep_parts;      // largest active molecule; CTAs per chain of the strict-mode sweep
    // reservoir pools (one per chain and active type): r_atom_base = first pool atom of residue r, r_cap = pool slots
    int has_reservoir, r_atoms, r_atom_base[MMC_MAX_RES], r_cap[MMC_MAX_RES];
    int shard_rank, shard_count;   // k/tile sharding of ONE system over several GPUs (1 handle per GPU); 0, 1 = off
};

// one pending trial move of one chain
struct Trial {
    int kind, res, mol, nat;
    int res2, nat2;        // NEW atom set: == res, nat except for a swap (the inserted species)
    int do_old, do_new;    // real-space sweeps wanted (ComputeOldEnergy / ComputeNewEnergy pair calls)
    int r_plus, r_minus;   // reciprocal update: A' = A + sum q phi(plus) - sum q phi(minus)
    int excl;              // 1: exclude slot (res, mol) from the sweep
    int accepted, error, draw;
    int rsv;               // creation with a reservoir: pool slot the geometry was copied from
    // Framework-energy cache (ChainState.fw_pair): fw_old = cached {LJ, Coulomb (reduced units)} of the selected molecule with the
    // framework at its current geometry, fw_old_ok = it was valid when the trial was built; after the sweep fw_new / fw_old hold
    // the sums of THIS trial when fw_new_ok / fw_old_fresh say so (the batched kernels only: fold_fw_cache), for commit_trial to store
    int fw_old_ok, fw_new_ok, fw_old_fresh;
    double fw_old[2], fw_new[2];
    double prob;
    double q[MMC_MAX_MOL_ATOMS];          // OLD atom set (the selected molecule)
    int type[MMC_MAX_MOL_ATOMS];
    double q2[MMC_MAX_MOL_ATOMS];         // NEW atom set
    int type2[MMC_MAX_MOL_ATOMS];
    double p_old[MMC_MAX_MOL_ATOMS][3];   // absolute positions, old geometry
    double p_new[MMC_MAX_MOL_ATOMS][3];   // absolute positions, proposed geometry
    double p_plus[MMC_MAX_MOL_ATOMS][3];  // positions whose phases are added (== p_new except deletion/reference)
    double com_new[3];
    double off_new[MMC_MAX_MOL_ATOMS][3];
    double old_e[MMC_NE], new_e[MMC_NE];
};

struct ChainState {
    int* count;        // [C][MMC_MAX_RES]
    double *gx, *gy, *gz;  // [C][g_atoms]
    double* gcom;      // [C][g_mols][3]
    double* goff;      // [C][g_atoms][3]
    double2* A;        // [C][2][nk]
    int* cur;          // [C] which A buffer is current
    double* energy;    // [C][MMC_NE]
    double* fug;       // [C][MMC_MAX_RES]
    double* temp;      // [C]
    double* tstep;     // [C]
    double* rstep;     // [C]
    mmc_counters* counters;  // [C]
    Trial* trial;      // [C]
    double* part;      // [C][sweep_parts][6]  (lj_old, coul_old, lj_new, coul_new, sum fW|A|^2, sum fW|A'|^2)
    long long* tape_pos;  // [C]
    int* rcount;          // [C][MMC_MAX_RES] molecules in the reservoir pool (reservoir%num_residues)
    double* roff;         // [C][r_atoms][3] their site offsets (reservoir%site_offset)
    unsigned int* chain_id;  // [C] Philox stream of each chain (default: the chain index); global ids make a sweep independent of how chains are dealt to GPUs
    unsigned long long* pair_count;  // [10] executed real-space work since mmc_init: distances formed, pairs inside reach; [2..9] = phase cycles of -DMMC_PHASE_TIMING builds
    int* error_flag;      // [1] MMC_ERRF_* bits, sticky until read by the host
    // [C][g_mols] framework part {LJ, Coulomb (reduced units)} of the pair energy of every guest molecule at its CURRENT geometry;
    // NaN = unknown.  The framework never moves, so the old-geometry framework sum of a translation / rotation / deletion is the
    // value summed when the molecule last moved: stored on accept, reused instead of re-summed (the reference re-sums it,
    // ComputeOldEnergy monte_carlo_utils.f90:347-396 -> energy_utils.f90:374-442).  Guest-guest terms are always re-summed.
    double2* fw_pair;
};

struct RngSrc {
    const double* tape;  // device pointer or nullptr
    long long tape_stride;
    unsigned long long seed;
    unsigned long long step;  // step index of this launch
};

// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int i = 0; i < 10; i++) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
        uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
// draw d of step s of chain c: same definition as oracle/maniac_oracle.c:orc_philox_uniform
__device__ __forceinline__ double philox_uniform(unsigned long long seed, uint32_t chain, unsigned long long step, uint32_t draw) {
    uint32_t c[4] = {(uint32_t)step, (uint32_t)(step >> 32), chain, draw >> 1};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    uint32_t lo = c[2 * (draw & 1)], hi = c[2 * (draw & 1) + 1];
    unsigned long long bits = (((unsigned long long)hi << 32) | lo) >> 11;
    return (double)bits * (1.0 / 9007199254740992.0);
}

// gfortran modulo(a, p), p > 0 (geometry_utils.f90:190,389): exact remainder in [0, p)
__device__ __forceinline__ double f_modulo(double a, double p, double invp) {
    double n = floor(a * invp);
    double r = fma(-n, p, a);
    if (r < 0.0) r += p;
    else if (r >= p) r -= p;
    return r;
}
// ComputeDistance geometry_utils.f90:385-393, one axis
__device__ __forceinline__ double min_image(double d, double L, double invL) {
    return f_modulo(d + 0.5 * L, L, invL) - 0.5 * L;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}
// deterministic block sum of NV values per thread, in two halves so that a caller can park one set of partial sums (rows
// row0 .. row0+NV-1 of smem[rows][NT/32]) and free its registers before it accumulates the next set
template <int NV, int NT>
__device__ __forceinline__ void block_sum_park(const double (&v)[NV], double* smem, int row0) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < NV; i++) {
        const double x = warp_sum(v[i]);
        if (lane == 0) smem[(row0 + i) * (NT / 32) + w] = x;
    }
}
// sums rows 0..NV-1 over the warps; result valid in thread 0.  Barrier before and after
template <int NV, int NT>
__device__ __forceinline__ void block_sum_gather(double (&v)[NV], const double* smem) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
    if (w == 0) {
#pragma unroll
        for (int i = 0; i < NV; i++) {
            double x = (lane < NT / 32) ? smem[i * (NT / 32) + lane] : 0.0;
            v[i] = warp_sum(x);
        }
    }
    __syncthreads();
}
// result valid in thread 0
template <int NV, int NT>
__device__ __forceinline__ void block_sum(double (&v)[NV], double* smem /* NV * NT/32 */) {
    block_sum_park<NV, NT>(v, smem, 0);
    block_sum_gather<NV, NT>(v, smem);
}

// ---- real-space Coulomb kernel g(s) = erfc(alpha sqrt(s)) / sqrt(s), s = r^2 ----------------------------
// Piecewise degree-MMC_GTAB_DEG polynomials on segments that follow the binary representation of s: 2^MMC_GTAB_MBITS segments
// per octave (exponent + top mantissa bits select the segment, the remaining mantissa bits give the local
// variable t in [-1, 1)), built at mmc_init for this system's alpha in long double (build_coulomb_table,
// maniac_gpu.cu).  Absolute error <= 5e-16 (it is at the rounding floor of the coefficients), i.e. a few
// 1e-16 relative where the terms matter (alpha r < 1.5) -- no sqrt, no exp, no division per pair.
// Reciprocal update of a translation as S_mol(k) (e^{i k d} - 1) (kernels_sweep.cuh, stage_terms): n + 1 product terms
// instead of 2 n.  On by default (A/B on one box: +1.5 %); -DMMC_NO_FACTORED_TRANSLATION restores the 2 n terms.
// Reciprocal sweep implementation (kernels_sweep.cuh): default = rank-n_term update dS(row, kz) = sum_t P_t(row) Z_t(kz) on
// the FP64 MMA unit (DMMA m8n8k4: the operands live in register fragments, so the per-term shared-memory loads of the
// item loop disappear); -DMMC_RECIP_ITEMS builds the round-1 item loop (one (kx, ky, +-kz) item per thread, P and Z from
// shared memory per term) for A/B runs.
#ifndef MMC_RECIP_ITEMS
#define MMC_RECIP_DMMA 1
#endif
#if !defined(MMC_NO_FACTORED_TRANSLATION) && !defined(MMC_RECIP_DMMA)
#define MMC_FACTORED_TRANSLATION 1
#endif
// Round 2: degree 8 on 16 segments per octave (was degree 9 on 8): measured against 40-digit values over r = 1.5 .. 37 A the
// relative error for r < 10 A drops from 6.1e-16 rms / 3.3e-15 max to 9e-17 rms / 2.7e-16 max (the rounding floor; glibc's
// erfc(x)/r, the CPU checker's kernel, sits at 3.2e-16 rms), with ONE coefficient load fewer per pair
// (tools/coulomb_table_accuracy.py); -DMMC_GTAB_DEG=9 -DMMC_GTAB_MBITS=3 rebuilds the round-1 table.
#ifndef MMC_GTAB_DEG
#define MMC_GTAB_DEG 8
#endif
#ifndef MMC_GTAB_MBITS
#define MMC_GTAB_MBITS 4
#endif
#define MMC_GTAB_MAXSEG (12 << MMC_GTAB_MBITS)   // 12 octaves (r = 1 A up to 64 A); also the COMPILE-TIME row stride of the layout
                                                 // [degree][segment]: the coefficient loads get immediate offsets (a run-time stride
                                                 // cost one integer multiply-add per load and 10 % of the real-space phase, r2b A/B)
__device__ __noinline__ double coul_kernel_slow(double s, double alpha) { const double r = sqrt(s); return erfc(alpha * r) / r; }
__device__ __forceinline__ double coul_kernel_tab(double s, const double* tab, int base, int i1, int i2) {
    const int hi = __double2hiint(s), lo = __double2loint(s);
    const int idx = (hi >> (20 - MMC_GTAB_MBITS)) - base;
    const double f = __hiloint2double(0x3FF00000 | ((hi & ((1 << (20 - MMC_GTAB_MBITS)) - 1)) << MMC_GTAB_MBITS) |
                                          (int)((unsigned)lo >> (32 - MMC_GTAB_MBITS)), lo << MMC_GTAB_MBITS);
    const double t = fma(2.0, f, -3.0);
    const double* c = tab + idx;
    // far segments are fitted with a lower degree (build_coulomb_table): their top one / two coefficients are not loaded
    double p = 0.0;
    if (idx < i1) p = c[MMC_GTAB_DEG * MMC_GTAB_MAXSEG];
    if (idx < i2) p = fma(p, t, c[(MMC_GTAB_DEG - 1) * MMC_GTAB_MAXSEG]);   // p == 0 here for [i1, i2): gives the coefficient itself
#pragma unroll
    for (int d = MMC_GTAB_DEG - 2; d >= 0; d--) p = fma(p, t, c[d * MMC_GTAB_MAXSEG]);
    return p;
}
// two evaluations in lockstep (old and new geometry of one pair): two independent Horner chains instead of two dependent
// ones back to back -- the real-space sweep is bound by the latency of these chains (r2c phase timing), not by a pipe
__device__ __forceinline__ void coul_kernel_tab2(double sa, double sb, const double* tab, int base, int i1, int i2, double& ga, double& gb) {
    const int hia = __double2hiint(sa), loa = __double2loint(sa), hib = __double2hiint(sb), lob = __double2loint(sb);
    const int ia = (hia >> (20 - MMC_GTAB_MBITS)) - base, ib = (hib >> (20 - MMC_GTAB_MBITS)) - base;
    const double fa = __hiloint2double(0x3FF00000 | ((hia & ((1 << (20 - MMC_GTAB_MBITS)) - 1)) << MMC_GTAB_MBITS) |
                                           (int)((unsigned)loa >> (32 - MMC_GTAB_MBITS)), loa << MMC_GTAB_MBITS);
    const double fb = __hiloint2double(0x3FF00000 | ((hib & ((1 << (20 - MMC_GTAB_MBITS)) - 1)) << MMC_GTAB_MBITS) |
                                           (int)((unsigned)lob >> (32 - MMC_GTAB_MBITS)), lob << MMC_GTAB_MBITS);
    const double ta = fma(2.0, fa, -3.0), tb = fma(2.0, fb, -3.0);
    const double* ca = tab + ia; const double* cb = tab + ib;
    double pa = 0.0, pb = 0.0;
    if (ia < i1) pa = ca[MMC_GTAB_DEG * MMC_GTAB_MAXSEG];
    if (ib < i1) pb = cb[MMC_GTAB_DEG * MMC_GTAB_MAXSEG];
    if (ia < i2) pa = fma(pa, ta, ca[(MMC_GTAB_DEG - 1) * MMC_GTAB_MAXSEG]);
    if (ib < i2) pb = fma(pb, tb, cb[(MMC_GTAB_DEG - 1) * MMC_GTAB_MAXSEG]);
#pragma unroll
    for (int d = MMC_GTAB_DEG - 2; d >= 0; d--) { pa = fma(pa, ta, ca[d * MMC_GTAB_MAXSEG]); pb = fma(pb, tb, cb[d * MMC_GTAB_MAXSEG]); }
    ga = pa; gb = pb;
}
// 1/a: MUFU.RCP64H seed + two Newton steps
__device__ __forceinline__ double mmc_rcp(double a) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
    double e = fma(-a, y, 1.0);
    y = fma(y, e, y);
    e = fma(-a, y, 1.0);
    return fma(y, e, y);
}
// minimum image by round-to-nearest: d - L * nint(d / L).  Same value as the reference's
// modulo(d + L/2, L) - L/2 (geometry_utils.f90:385-393) up to rounding; the 1.5*2^52 constant rounds to integer
__device__ __forceinline__ double min_image_fast(double d, double L, double invL) {
    const double big = 6755399441055744.0;
    const double n = fma(d, invL, big) - big;
    return fma(-n, L, d);
}

// Triclinic nearest image of a separation (ComputeDistance's triclinic branch, geometry_utils.f90:397-411, with the lattice
// vectors a, b, c as the shifts).  The cascade z -> y -> x lands in the brick [-l/2, l/2]^3, a fundamental domain of the
// lattice; a brick point shorter than half the shortest edge IS the nearest image, any other one gets the reference's
// 27-image search around it.
__device__ __forceinline__ void min_image_tri_brick(const DevConst& dc, double& dx, double& dy, double& dz) {
    const double big = 6755399441055744.0;
    double n = fma(dz, dc.invL[2], big) - big;
    dz = fma(-n, dc.L[2], dz); dy = fma(-n, dc.tilt[2], dy); dx = fma(-n, dc.tilt[1], dx);
    n = fma(dy, dc.invL[1], big) - big;
    dy = fma(-n, dc.L[1], dy); dx = fma(-n, dc.tilt[0], dx);
    n = fma(dx, dc.invL[0], big) - big;
    dx = fma(-n, dc.L[0], dx);
}
// out of line and fed plain values: a reference to the kernel-parameter DevConst here would force a local copy of it
__device__ __noinline__ double3 min_image_tri_search(double lx, double ly, double lz, double xy, double xz, double yz,
                                                     double x0, double y0, double z0) {
    double3 b = make_double3(x0, y0, z0);
    double best = x0 * x0 + y0 * y0 + z0 * z0;
    for (int sa = -1; sa <= 1; sa++)
        for (int sb = -1; sb <= 1; sb++)
            for (int sc = -1; sc <= 1; sc++) {
                const double x = x0 + sa * lx + sb * xy + sc * xz;
                const double y = y0 + sb * ly + sc * yz;
                const double z = z0 + sc * lz;
                const double r2 = x * x + y * y + z * z;
                if (r2 < best) { best = r2; b = make_double3(x, y, z); }
            }
    return b;
}
__device__ __forceinline__ void min_image_tri(const DevConst& dc, double& dx, double& dy, double& dz) {
    min_image_tri_brick(dc, dx, dy, dz);
    if (dx * dx + dy * dy + dz * dz >= dc.half_min2) {
        const double3 b = min_image_tri_search(dc.L[0], dc.L[1], dc.L[2], dc.tilt[0], dc.tilt[1], dc.tilt[2], dx, dy, dz);
        dx = b.x; dy = b.y; dz = b.z;
    }
}
// separation under the cell's periodicity, exact form (ComputeDistance geometry_utils.f90:385-411)
__device__ __forceinline__ void min_image3(const DevConst& dc, double& dx, double& dy, double& dz) {
    if (dc.tri) { min_image_tri(dc, dx, dy, dz); return; }
    dx = min_image(dx, dc.L[0], dc.invL[0]); dy = min_image(dy, dc.L[1], dc.invL[1]); dz = min_image(dz, dc.L[2], dc.invL[2]);
}
// coordinate `ax` of (a b c)^-1 p: the argument of ComputeAtomPhase (ewald_phase.f90:41-63) is 2 pi times this
__device__ __forceinline__ double frac_coord(const DevConst& dc, const double* p, int ax) {
    if (!dc.tri) return dc.invL[ax] * p[ax];
    const double sz = dc.invL[2] * p[2];
    if (ax == 2) return sz;
    const double sy = dc.invL[1] * (p[1] - dc.tilt[2] * sz);
    if (ax == 1) return sy;
    return dc.invL[0] * (p[0] - dc.tilt[0] * sy - dc.tilt[1] * sz);
}

// LookupTabulated tabulated_utils.f90:87-113
__device__ __forceinline__ double table_lookup(const double* f, double dx, double r) {
    const int n = 5000;  // TABULATED_POINTS parameters.f90:41
    if (r <= 0.0) return f[0];
    if (r >= n * dx) return 0.0;
    int i = (int)(r / dx);
    double f1 = f[i], f2 = f[i + 1];
    double t = (r - i * dx) / dx;
    return (1.0 - t) * f1 + t * f2;
}
