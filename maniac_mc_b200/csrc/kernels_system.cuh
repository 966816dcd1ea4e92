// kernels_system.cuh -- single-point energy ComputeSystemEnergy (energy_utils.f90:18-35) on the device:
// total real-space pair energy (i<j, intra-molecular pairs excluded, Q13), full structure factor
// S(k) + reciprocal energy, Ewald self term and intra-residue correction.  Runs once per chain at
// start-up (and in tests as the consistency check of the tracked energies).
#pragma once
#include "common.cuh"

#define SYS_THREADS 256
#define SK_CHUNK 32

struct SysScratch {
    double *x, *y, *z, *q;   // [N] flat atoms of one chain: framework then guest slots
    int *type, *mol;         // mol = -1 for an empty guest slot
    int *tile_min, *tile_max;  // molecule-id range per tile of SYS_THREADS atoms
    int n, n_tiles;
    double* part;            // partial sums
    double2* sk_part;        // [n_super][nk]
    int n_super;
    double* result;          // [8]
};

__global__ void k_sys_gather(DevConst dc, ChainState cs, SysScratch ss, int c, int n_fw_mols) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= ss.n) return;
    int mol = -1;
    if (j < dc.n_fw) {
        ss.x[j] = dc.fx[j]; ss.y[j] = dc.fy[j]; ss.z[j] = dc.fz[j]; ss.q[j] = dc.fq[j]; ss.type[j] = dc.ftype[j];
        mol = dc.fmol[j];
    } else {
        const int g = j - dc.n_fw;
        int r = 0;
        for (int i = 0; i < dc.n_active; i++) { const int rr = dc.active_list[i]; if (g >= dc.g_atom_base[rr]) r = rr; }
        const int na = dc.natoms[r];
        const int m = (g - dc.g_atom_base[r]) / na, a = (g - dc.g_atom_base[r]) % na;
        const size_t ia = (size_t)c * dc.g_atoms + g;
        ss.x[j] = cs.gx[ia]; ss.y[j] = cs.gy[ia]; ss.z[j] = cs.gz[ia]; ss.q[j] = dc.q[r][a]; ss.type[j] = dc.type[r][a];
        if (m < cs.count[c * MMC_MAX_RES + r]) mol = n_fw_mols + dc.g_mol_base[r] + m;
    }
    ss.mol[j] = mol;
    if (mol >= 0) { atomicMin(&ss.tile_min[j / SYS_THREADS], mol); atomicMax(&ss.tile_max[j / SYS_THREADS], mol); }
}

// LennardJonesEnergy energy_utils.f90:192-218 and CoulombEnergy :225-256 (with the optional table branch)
__device__ __forceinline__ void sys_pair_eval(const DevConst& dc, double r, double2 lj, double qi, double qj, double& e_lj, double& e_c) {
    if (r < dc.rc) {
        double r6, r12;
        if (dc.use_table) {
            const double s2 = lj.x * lj.x, s6 = s2 * s2 * s2;
            r6 = s6 / table_lookup(dc.tab_r6, dc.tab_dx, r);
            r12 = s6 * s6 / table_lookup(dc.tab_r12, dc.tab_dx, r);
        } else { const double sr = lj.x / r, sr2 = sr * sr; r6 = sr2 * sr2 * sr2; r12 = r6 * r6; }
        e_lj += lj.y * (r12 - r6);
    }
    if (fabs(qi) >= MMC_ERR && fabs(qj) >= MMC_ERR && r >= MMC_ERR) {
        if (dc.use_table) e_c += qi * qj * table_lookup(dc.tab_erfc, dc.tab_dx, r);
        else e_c += qi * qj * erfc(dc.alpha * r) / r;
    }
}

// grid = (n_tiles, n_tiles); block (bi, bj) with bj >= bi evaluates tile_i x tile_j, pair (i, j) kept if j > i
__global__ void __launch_bounds__(SYS_THREADS) k_sys_pair(DevConst dc, SysScratch ss) {
    __shared__ double sx[SYS_THREADS], sy[SYS_THREADS], sz[SYS_THREADS], sq[SYS_THREADS];
    __shared__ int st[SYS_THREADS], smol[SYS_THREADS];
    __shared__ double red[2 * SYS_THREADS / 32];
    const int bi = blockIdx.x, bj = blockIdx.y;
    double acc[2] = {0.0, 0.0};
    double* out = ss.part + ((size_t)bi * ss.n_tiles + bj) * 2;
    bool skip = bj < bi;
    if (!skip) {
        const int a0 = ss.tile_min[bi], a1 = ss.tile_max[bi], b0 = ss.tile_min[bj], b1 = ss.tile_max[bj];
        if (a1 < 0 || b1 < 0) skip = true;                          // a tile of empty slots
        else if (a0 == a1 && b0 == b1 && a0 == b0) skip = true;     // both inside one molecule (framework)
    }
    if (skip) { if (threadIdx.x == 0) { out[0] = 0.0; out[1] = 0.0; } return; }
    const int j0 = bj * SYS_THREADS + threadIdx.x;
    if (j0 < ss.n) { sx[threadIdx.x] = ss.x[j0]; sy[threadIdx.x] = ss.y[j0]; sz[threadIdx.x] = ss.z[j0]; sq[threadIdx.x] = ss.q[j0];
                     st[threadIdx.x] = ss.type[j0]; smol[threadIdx.x] = ss.mol[j0]; }
    else smol[threadIdx.x] = -1;
    __syncthreads();
    const int i = bi * SYS_THREADS + threadIdx.x;
    if (i < ss.n && ss.mol[i] >= 0) {
        const double xi = ss.x[i], yi = ss.y[i], zi = ss.z[i], qi = ss.q[i];
        const int ti = ss.type[i], mi = ss.mol[i];
        for (int jj = 0; jj < SYS_THREADS; jj++) {
            const int j = bj * SYS_THREADS + jj;
            if (j <= i || smol[jj] < 0 || smol[jj] == mi) continue;
            double dx = sx[jj] - xi, dy = sy[jj] - yi, dz = sz[jj] - zi;
            min_image3(dc, dx, dy, dz);
            const double r = sqrt(dx * dx + dy * dy + dz * dz);
            sys_pair_eval(dc, r, dc.lj[ti * dc.ntypes + st[jj]], qi, sq[jj], acc[0], acc[1]);
        }
    }
    block_sum<2, SYS_THREADS>(acc, red);
    if (threadIdx.x == 0) { out[0] = acc[0]; out[1] = acc[1]; }
}

// deterministic sum of n rows of `width` doubles -> out[width]; one block
__global__ void __launch_bounds__(SYS_THREADS) k_sum_rows(const double* part, long long n, int width, double* out, double scale0, double scale1) {
    __shared__ double red[2 * SYS_THREADS / 32];
    double v[2] = {0.0, 0.0};
    for (long long i = threadIdx.x; i < n; i += SYS_THREADS) { v[0] += part[i * width]; if (width > 1) v[1] += part[i * width + 1]; }
    block_sum<2, SYS_THREADS>(v, red);
    if (threadIdx.x == 0) { out[0] = v[0] * scale0; if (width > 1) out[1] = v[1] * scale1; }
}

// Full structure factor (ComputeAllFourierTerms ewald_phase.f90:340 + ComputeRecipAmplitude ewald_energy.f90:40):
// grid = (n_k_tiles, n_super); each block walks its share of atom chunks, builds the chunk's 1-D phase
// tables in shared memory (direct cos/sin(k*theta) as ComputePhaseFactors1D) and accumulates
// sum_atoms q * phix(kx) phiy(ky) phiz(kz) for its k-vectors.
__global__ void __launch_bounds__(RECIP_THREADS) k_sys_sk(DevConst dc, SysScratch ss, int chunk) {
    extern __shared__ double2 s_tab[];  // [chunk][3][kstride]
    const int ks = dc.kstride;
    const int ik = blockIdx.x * RECIP_THREADS + threadIdx.x;
    int4 kv = make_int4(0, 0, 0, 0);
    if (ik < dc.nk) kv = dc.kvec[ik];
    double are = 0.0, aim = 0.0;
    const int n_chunks = (ss.n + chunk - 1) / chunk;
    const int per = (n_chunks + gridDim.y - 1) / gridDim.y;
    for (int ch = blockIdx.y * per; ch < min(n_chunks, (int)(blockIdx.y + 1) * per); ch++) {
        __syncthreads();
        for (int i = threadIdx.x; i < chunk * 3 * ks; i += blockDim.x) {
            const int k = i % ks, ax = (i / ks) % 3, a = i / (3 * ks);
            const int j = ch * chunk + a;
            double2 v = make_double2(0.0, 0.0);
            if (j < ss.n && k <= dc.kmax[ax] && ss.mol[j] >= 0) {
                const double pos[3] = {ss.x[j], ss.y[j], ss.z[j]};
                const double theta = MMC_TWOPI * frac_coord(dc, pos, ax);
                double s, co; sincos((double)k * theta, &s, &co);
                v = make_double2(co, s);
            }
            s_tab[i] = v;
        }
        __syncthreads();
        if (ik < dc.nk) {
            for (int a = 0; a < chunk; a++) {
                const int j = ch * chunk + a;
                if (j >= ss.n) break;
                if (ss.mol[j] < 0) continue;
                const double2* tb = s_tab + (size_t)a * 3 * ks;
                double2 px = tb[kv.x];
                double2 py = tb[ks + abs(kv.y)]; if (kv.y < 0) py.y = -py.y;
                double2 pz = tb[2 * ks + abs(kv.z)]; if (kv.z < 0) pz.y = -pz.y;
                const double2 pxy = make_double2(px.x * py.x - px.y * py.y, px.x * py.y + px.y * py.x);
                const double2 p = make_double2(pxy.x * pz.x - pxy.y * pz.y, pxy.x * pz.y + pxy.y * pz.x);
                const double q = ss.q[j];
                are += q * p.x; aim += q * p.y;
            }
        }
    }
    if (ik < dc.nk) ss.sk_part[(size_t)blockIdx.y * dc.nk + ik] = make_double2(are, aim);
}

// sums the super-chunk partials in order -> S(k); energy partial per block (ComputeReciprocalEnergy
// ewald_energy.f90:105-148); stores S(k) as the tracked A(k) (init_sk=full) or zeros it (Q1 default)
__global__ void __launch_bounds__(RECIP_THREADS) k_sys_sk_reduce(DevConst dc, ChainState cs, SysScratch ss, int c, int store) {
    __shared__ double red[2 * RECIP_THREADS / 32];
    const int ik = blockIdx.x * RECIP_THREADS + threadIdx.x;
    double v[2] = {0.0, 0.0};
    if (ik < dc.nk) {
        double are = 0.0, aim = 0.0;
        for (int s = 0; s < ss.n_super; s++) { const double2 p = ss.sk_part[(size_t)s * dc.nk + ik]; are += p.x; aim += p.y; }
        v[0] = dc.fw[ik] * (are * are + aim * aim);
        double2* A = cs.A + ((size_t)c * 2 + cs.cur[c]) * dc.nk;
        A[ik] = store ? make_double2(are, aim) : make_double2(0.0, 0.0);
    }
    block_sum<2, RECIP_THREADS>(v, red);
    if (threadIdx.x == 0) { ss.part[blockIdx.x * 2] = v[0]; ss.part[blockIdx.x * 2 + 1] = 0.0; }
}

// ComputeEwaldSelf energy_utils.f90:307-330 (framework part: -alpha/sqrt(pi) sum q^2 over all inactive
// atoms) and ComputeTotalIntraResidueCoulombEnergy :55-81 (active residues only).  One block.
__global__ void __launch_bounds__(SYS_THREADS) k_sys_self_intra(DevConst dc, ChainState cs, int c, double* out /* [2] */) {
    __shared__ double red[2 * SYS_THREADS / 32];
    double v[2] = {0.0, 0.0};
    const double f = dc.alpha / sqrt(MMC_PI);
    for (int j = threadIdx.x; j < dc.n_fw; j += SYS_THREADS) { const double q = dc.fq[j]; if (fabs(q) >= MMC_ERR) v[0] -= f * (q * q); }
    for (int i = 0; i < dc.n_active; i++) {
        const int r = dc.active_list[i], nat = dc.natoms[r], cnt = cs.count[c * MMC_MAX_RES + r];
        for (int m = threadIdx.x; m < cnt; m += SYS_THREADS) {
            double p[MMC_MAX_MOL_ATOMS][3];
            for (int a = 0; a < nat; a++) {
                const size_t ia = (size_t)c * dc.g_atoms + dc.g_atom_base[r] + (size_t)m * nat + a;
                p[a][0] = cs.gx[ia]; p[a][1] = cs.gy[ia]; p[a][2] = cs.gz[ia];
            }
            double e = 0.0;
            for (int a1 = 0; a1 < nat - 1; a1++)
                for (int a2 = a1 + 1; a2 < nat; a2++) {
                    double dx = p[a2][0] - p[a1][0], dy = p[a2][1] - p[a1][1], dz = p[a2][2] - p[a1][2];
                    min_image3(dc, dx, dy, dz);
                    const double rr = sqrt(dx * dx + dy * dy + dz * dz);
                    if (rr > MMC_ERR) e += dc.q[r][a1] * dc.q[r][a2] * (erfc(dc.alpha * rr) - 1.0) / rr;
                }
            v[1] += e * dc.coul_const;
        }
    }
    block_sum<2, SYS_THREADS>(v, red);
    if (threadIdx.x == 0) {
        double self = v[0] * dc.coul_const;
        for (int i = 0; i < dc.n_active; i++) { const int r = dc.active_list[i]; self += dc.self_e[r] * cs.count[c * MMC_MAX_RES + r]; }
        out[0] = self; out[1] = v[1];
    }
}

// W(k) = exp(-k^2 / 4 alpha^2) / k^2 (ComputeReciprocalWeights ewald_kvectors.f90:225-246) times the
// symmetry form factor (ComputeSymmetryFormFactor :167); k^2 from ComputeCartesianKSquared :134
__global__ void k_init_weights(DevConst dc, double* fw, double* w_only) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= dc.nk) return;
    const int4 kv = dc.kvec[i];
    double vx = (double)kv.x * (MMC_TWOPI * dc.invL[0]), vy = (double)kv.y * (MMC_TWOPI * dc.invL[1]),
           vz = (double)kv.z * (MMC_TWOPI * dc.invL[2]);
    if (dc.tri) {   // k = 2 pi (a b c)^-T n
        const double gx = (double)kv.x * dc.invL[0];
        const double gy = ((double)kv.y - dc.tilt[0] * gx) * dc.invL[1];
        const double gz = ((double)kv.z - dc.tilt[1] * gx - dc.tilt[2] * gy) * dc.invL[2];
        vx = MMC_TWOPI * gx; vy = MMC_TWOPI * gy; vz = MMC_TWOPI * gz;
    }
    const double k2 = vx * vx + vy * vy + vz * vz;
    const double w = exp(-k2 / (4.0 * (dc.alpha * dc.alpha))) / k2;
    w_only[i] = w;
    fw[i] = (kv.x == 0 ? 1.0 : 2.0) * w;
}

// tabulated_utils.f90:19-83 : tables of erfc(alpha r)/r, r^6, r^12 on [0, rc], 5001 points
__global__ void k_init_tables(double alpha, double rc, double* t_erfc, double* t_r6, double* t_r12) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > 5000) return;
    const double dx = rc / 5000.0, r = i * dx;
    if (r < MMC_ERR) { t_erfc[i] = 2.0 * alpha / sqrt(MMC_PI); t_r6[i] = 0.0; t_r12[i] = 0.0; }
    else {
        t_erfc[i] = erfc(alpha * r) / r;
        const double r2 = r * r, r6 = r2 * r2 * r2;
        t_r6[i] = r6; t_r12[i] = r6 * r6;
    }
}
__global__ void k_table_lookup(DevConst dc, int which, const double* r, int n, double* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = table_lookup(which == 0 ? dc.tab_erfc : which == 1 ? dc.tab_r6 : dc.tab_r12, dc.tab_dx, r[i]);
}
// erfc(x) as the per-move sweep evaluates it: r = x / alpha, erfc = r * g(r^2) with the Coulomb kernel table
// (slow path outside the table's domain)
__global__ void k_erfc_test(DevConst dc, const double* x, int n, double* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double r = x[i] / dc.alpha, s = r * r;
    const double g = (s >= dc.gtab_smin && s < dc.gtab_smax) ? coul_kernel_tab(s, dc.gtab, dc.gtab_base, dc.gtab_deg_start[0], dc.gtab_deg_start[1]) : coul_kernel_slow(s, dc.alpha);
    out[i] = r * g;
}

// FP64 peak: 8 independent DFMA chains per thread, register resident
__global__ void k_fp64_peak(double* out, int iters) {
    double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double m = 0.999999, b = 1e-7;
    for (int i = 0; i < iters; i++) {
        a0 = fma(a0, m, b); a1 = fma(a1, m, b); a2 = fma(a2, m, b); a3 = fma(a3, m, b);
        a4 = fma(a4, m, b); a5 = fma(a5, m, b); a6 = fma(a6, m, b); a7 = fma(a7, m, b);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

// ====================================================================================================================
// Start-up for MANY chains at once (mmc_system_energy_all).  ComputeSystemEnergy (energy_utils.f90:18-35) splits into a part
// that is the same for every chain of a handle -- the inactive residues ("framework"): their structure factor
// S_fw(k) = sum_fw q e^{i k r} (ComputeRecipAmplitude over the inactive atoms, ewald_energy.f90:40-74), their mutual pair
// energy and their Ewald self term (energy_utils.f90:307-330) -- computed ONCE per handle, and a per-chain part over the
// guests only: guest-framework and guest-guest pairs, S_guests(k), self and intra-residue terms.  A_chain = S_fw + S_guests.
// The single-chain path above (mmc_system_energy) is kept as the independent evaluation the tests compare against.

// sums the super-chunk partials of k_sys_sk in order -> out[k] (the framework's S_fw)
__global__ void __launch_bounds__(RECIP_THREADS) k_sys_sk_store(DevConst dc, SysScratch ss, double2* out) {
    const int ik = blockIdx.x * RECIP_THREADS + threadIdx.x;
    if (ik >= dc.nk) return;
    double are = 0.0, aim = 0.0;
    for (int s = 0; s < ss.n_super; s++) { const double2 p = ss.sk_part[(size_t)s * dc.nk + ik]; are += p.x; aim += p.y; }
    out[ik] = make_double2(are, aim);
}

// -alpha/sqrt(pi) sum q^2 over the inactive atoms (ComputeEwaldSelf, framework share), in K; one block
__global__ void __launch_bounds__(SYS_THREADS) k_sys_fw_self(DevConst dc, double* out) {
    __shared__ double red[2 * SYS_THREADS / 32];
    double v[2] = {0.0, 0.0};
    const double f = dc.alpha / sqrt(MMC_PI);
    for (int j = threadIdx.x; j < dc.n_fw; j += SYS_THREADS) { const double q = dc.fq[j]; if (fabs(q) >= MMC_ERR) v[0] -= f * (q * q); }
    block_sum<2, SYS_THREADS>(v, red);
    if (threadIdx.x == 0) out[0] = v[0] * dc.coul_const;
}

struct SysAll {
    const double2* sfw;    // [nk] framework structure factor
    const double* fwc;     // [3] framework-framework LJ, Coulomb (K), framework self (K)
    double* pair_part;     // [C][n_other_tiles][2]
    double* recip_part;    // [C][n_k_tiles]
    double* out;           // [C][MMC_NE] (device copy of the results, same values as cs.energy)
    int n_other_tiles, n_fw_mols, store_sk, chain0;
};

// guest slot g of a chain -> (residue, molecule, atom)
__device__ __forceinline__ void guest_slot(const DevConst& dc, int g, int& r, int& m, int& a) {
    r = dc.active_list[0];
    for (int i = 0; i < dc.n_active; i++) { const int rr = dc.active_list[i]; if (g >= dc.g_atom_base[rr]) r = rr; }
    const int na = dc.natoms[r], l = g - dc.g_atom_base[r];
    m = l / na; a = l - m * na;
}

// grid = (n_other_tiles, chains): block (T, c) holds 256 "other" atoms j of the flat order framework ++ guest slots of chain
// c (one per thread) and walks the chain's guest atoms i; the unordered inter-molecular pair (i, j) is taken where j is a
// framework atom or a guest atom behind i in the flat order (the ordering test of energy_utils.f90:157-158)
__global__ void __launch_bounds__(SYS_THREADS) k_sys_pair_guests(DevConst dc, ChainState cs, SysAll sa) {
    __shared__ double sx[SYS_THREADS], sy[SYS_THREADS], sz[SYS_THREADS], sq[SYS_THREADS];
    __shared__ int st[SYS_THREADS], smol[SYS_THREADS], sflat[SYS_THREADS];
    __shared__ double red[2 * SYS_THREADS / 32];
    const int c = sa.chain0 + blockIdx.y, T = blockIdx.x, tid = threadIdx.x;
    const int j = T * SYS_THREADS + tid;
    double xj = 0, yj = 0, zj = 0, qj = 0; int tj = 0, mj = -1;
    if (dc.n_active > 0) {
        if (j < dc.n_fw) { xj = dc.fx[j]; yj = dc.fy[j]; zj = dc.fz[j]; qj = dc.fq[j]; tj = dc.ftype[j]; mj = dc.fmol[j]; }
        else if (j < dc.n_fw + dc.g_atoms) {
            int r, m, a; guest_slot(dc, j - dc.n_fw, r, m, a);
            if (m < cs.count[c * MMC_MAX_RES + r]) {
                const size_t ia = (size_t)c * dc.g_atoms + (j - dc.n_fw);
                xj = cs.gx[ia]; yj = cs.gy[ia]; zj = cs.gz[ia]; qj = dc.q[r][a]; tj = dc.type[r][a];
                mj = sa.n_fw_mols + dc.g_mol_base[r] + m;
            }
        }
    }
    double acc[2] = {0.0, 0.0};
    for (int ia = 0; ia < dc.n_active; ia++) {
        const int r = dc.active_list[ia], na = dc.natoms[r];
        const int n_at = cs.count[c * MMC_MAX_RES + r] * na;
        for (int i0 = 0; i0 < n_at; i0 += SYS_THREADS) {
            __syncthreads();
            const int i = i0 + tid;
            if (i < n_at) {
                const int m = i / na, a = i - m * na;
                const size_t gi = (size_t)c * dc.g_atoms + dc.g_atom_base[r] + i;
                sx[tid] = cs.gx[gi]; sy[tid] = cs.gy[gi]; sz[tid] = cs.gz[gi]; sq[tid] = dc.q[r][a]; st[tid] = dc.type[r][a];
                smol[tid] = sa.n_fw_mols + dc.g_mol_base[r] + m; sflat[tid] = dc.n_fw + dc.g_atom_base[r] + i;
            }
            __syncthreads();
            if (mj >= 0) {
                const int nn = min(SYS_THREADS, n_at - i0);
                for (int ii = 0; ii < nn; ii++) {
                    if (smol[ii] == mj || !(j < dc.n_fw || j > sflat[ii])) continue;
                    // ComputeDistance: delta = pos(second) - pos(first), first = the lower flat index
                    const bool i_first = sflat[ii] < j;
                    double dx = i_first ? xj - sx[ii] : sx[ii] - xj, dy = i_first ? yj - sy[ii] : sy[ii] - yj, dz = i_first ? zj - sz[ii] : sz[ii] - zj;
                    min_image3(dc, dx, dy, dz);
                    const double rr = sqrt(dx * dx + dy * dy + dz * dz);
                    const int ti = i_first ? st[ii] : tj, t2 = i_first ? tj : st[ii];
                    sys_pair_eval(dc, rr, dc.lj[ti * dc.ntypes + t2], i_first ? sq[ii] : qj, i_first ? qj : sq[ii], acc[0], acc[1]);
                }
            }
        }
    }
    __syncthreads();
    block_sum<2, SYS_THREADS>(acc, red);
    if (tid == 0) { double* o = sa.pair_part + ((size_t)blockIdx.y * sa.n_other_tiles + T) * 2; o[0] = acc[0]; o[1] = acc[1]; }
}

// grid = (n_k_tiles, chains): S_guests(k) of chain c for 256 k-vectors (1-D phase tables of a chunk of guest atoms in shared
// memory, direct cos / sin as ComputePhaseFactors1D), A = S_fw + S_guests stored as the tracked A(k) (init_sk = full) or zeroed
// (Q1 default), energy partial sum f W |A|^2
__global__ void __launch_bounds__(RECIP_THREADS) k_sys_sk_guests(DevConst dc, ChainState cs, SysAll sa, int chunk) {
    extern __shared__ double2 s_tab[];  // [chunk][3][kstride]
    __shared__ double s_q[SK_CHUNK];
    __shared__ double red[2 * RECIP_THREADS / 32];
    const int ks = dc.kstride, c = sa.chain0 + blockIdx.y;
    const int ik = blockIdx.x * RECIP_THREADS + threadIdx.x;
    int4 kv = make_int4(0, 0, 0, 0);
    if (ik < dc.nk) kv = dc.kvec[ik];
    double are = 0.0, aim = 0.0;
    for (int ia = 0; ia < dc.n_active; ia++) {
        const int r = dc.active_list[ia], na = dc.natoms[r];
        const int n_at = cs.count[c * MMC_MAX_RES + r] * na;
        const size_t base = (size_t)c * dc.g_atoms + dc.g_atom_base[r];
        for (int a0 = 0; a0 < n_at; a0 += chunk) {
            const int nn = min(chunk, n_at - a0);
            __syncthreads();
            for (int i = threadIdx.x; i < nn * 3 * ks; i += blockDim.x) {
                const int k = i % ks, ax = (i / ks) % 3, a = i / (3 * ks);
                double2 v = make_double2(0.0, 0.0);
                if (k <= dc.kmax[ax]) {
                    const double pos[3] = {cs.gx[base + a0 + a], cs.gy[base + a0 + a], cs.gz[base + a0 + a]};
                    const double theta = MMC_TWOPI * frac_coord(dc, pos, ax);
                    double s, co; sincos((double)k * theta, &s, &co);
                    v = make_double2(co, s);
                }
                s_tab[i] = v;
            }
            if (threadIdx.x < nn) s_q[threadIdx.x] = dc.q[r][(a0 + threadIdx.x) % na];
            __syncthreads();
            if (ik < dc.nk) {
                for (int a = 0; a < nn; a++) {
                    const double2* tb = s_tab + (size_t)a * 3 * ks;
                    double2 px = tb[kv.x];
                    double2 py = tb[ks + abs(kv.y)]; if (kv.y < 0) py.y = -py.y;
                    double2 pz = tb[2 * ks + abs(kv.z)]; if (kv.z < 0) pz.y = -pz.y;
                    const double2 pxy = make_double2(px.x * py.x - px.y * py.y, px.x * py.y + px.y * py.x);
                    const double2 p = make_double2(pxy.x * pz.x - pxy.y * pz.y, pxy.x * pz.y + pxy.y * pz.x);
                    const double q = s_q[a];
                    are += q * p.x; aim += q * p.y;
                }
            }
        }
    }
    double v[2] = {0.0, 0.0};
    if (ik < dc.nk) {
        const double2 f = sa.sfw[ik];
        are += f.x; aim += f.y;
        v[0] = dc.fw[ik] * (are * are + aim * aim);
        double2* A = cs.A + ((size_t)c * 2 + cs.cur[c]) * dc.nk;
        A[ik] = sa.store_sk ? make_double2(are, aim) : make_double2(0.0, 0.0);
    }
    __syncthreads();
    block_sum<2, RECIP_THREADS>(v, red);
    if (threadIdx.x == 0) sa.recip_part[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = v[0];
}

// grid = chains: ordered sums of the chain's partials, guest self and intra terms, tracked energies
__global__ void __launch_bounds__(SYS_THREADS) k_sys_finish_all(DevConst dc, ChainState cs, SysAll sa, int n_k_tiles) {
    __shared__ double red[4 * SYS_THREADS / 32];
    const int c = sa.chain0 + blockIdx.x;
    double v[4] = {0.0, 0.0, 0.0, 0.0};   // lj, coul, recip, intra
    const double* pp = sa.pair_part + (size_t)blockIdx.x * sa.n_other_tiles * 2;
    for (int i = threadIdx.x; i < sa.n_other_tiles; i += SYS_THREADS) { v[0] += pp[2 * i]; v[1] += pp[2 * i + 1]; }
    const double* rp = sa.recip_part + (size_t)blockIdx.x * n_k_tiles;
    for (int i = threadIdx.x; i < n_k_tiles; i += SYS_THREADS) v[2] += rp[i];
    for (int i = 0; i < dc.n_active; i++) {   // ComputeTotalIntraResidueCoulombEnergy energy_utils.f90:55-81 (active residues only)
        const int r = dc.active_list[i], nat = dc.natoms[r], cnt = cs.count[c * MMC_MAX_RES + r];
        for (int m = threadIdx.x; m < cnt; m += SYS_THREADS) {
            double e = 0.0;
            const size_t b = (size_t)c * dc.g_atoms + dc.g_atom_base[r] + (size_t)m * nat;
            for (int a1 = 0; a1 < nat - 1; a1++)
                for (int a2 = a1 + 1; a2 < nat; a2++) {
                    double dx = cs.gx[b + a2] - cs.gx[b + a1], dy = cs.gy[b + a2] - cs.gy[b + a1], dz = cs.gz[b + a2] - cs.gz[b + a1];
                    min_image3(dc, dx, dy, dz);
                    const double rr = sqrt(dx * dx + dy * dy + dz * dz);
                    if (rr > MMC_ERR) e += dc.q[r][a1] * dc.q[r][a2] * (erfc(dc.alpha * rr) - 1.0) / rr;
                }
            v[3] += e * dc.coul_const;
        }
    }
    block_sum<4, SYS_THREADS>(v, red);
    if (threadIdx.x == 0) {
        double self = sa.fwc[2];
        for (int i = 0; i < dc.n_active; i++) { const int r = dc.active_list[i]; self += dc.self_e[r] * cs.count[c * MMC_MAX_RES + r]; }
        double* E = cs.energy + (size_t)c * MMC_NE;
        E[MMC_E_NONCOUL] = sa.fwc[0] + v[0];
        E[MMC_E_COUL] = sa.fwc[1] + v[1] * dc.coul_const;
        E[MMC_E_RECIP] = v[2] * dc.recip_const;
        E[MMC_E_SELF] = self; E[MMC_E_INTRA] = v[3];
        E[MMC_E_TOTAL] = E[MMC_E_RECIP] + E[MMC_E_NONCOUL] + E[MMC_E_COUL] + E[MMC_E_SELF] + E[MMC_E_INTRA];   // energy_utils.f90:32-33
        if (sa.out) for (int i = 0; i < MMC_NE; i++) sa.out[(size_t)blockIdx.x * MMC_NE + i] = E[i];
    }
}
