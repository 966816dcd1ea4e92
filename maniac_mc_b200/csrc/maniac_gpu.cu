// maniac_gpu.cu -- handle management and the C ABI of libmaniacgpu.so (include/maniac_gpu.h).
// Host code here only packs descriptions, owns device buffers, and enqueues kernels; every energy
// is computed on the device.  There is no CPU fallback.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "common.cuh"
#include "kernels_move.cuh"
#include "kernels_sweep.cuh"
#include "kernels_chain.cuh"
#include "kernels_system.cuh"
#include <algorithm>
#include <atomic>
#include <chrono>
#include <mutex>
#include <numeric>

static thread_local std::string g_err;
static int fail(const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    g_err = buf;
    return 1;
}
#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)
#define CHECK_CHAIN(h, c) do { if (!(h)) return fail("null handle"); if ((c) < 0 || (c) >= (h)->dc.n_chains) return fail("chain %d out of range [0,%d)", (c), (h)->dc.n_chains); } while (0)

struct mmc_handle {
    DevConst dc;
    ChainState cs;
    SysScratch ss;
    cudaStream_t stream = nullptr;
    int device = 0;
    int init_sk = 0;
    int n_fw_mols = 0;
    int sk_chunk = SK_CHUNK;
    size_t sk_smem = 0;
    size_t sweep_smem = 0, chain_smem = 0;  // dynamic shared memory of k_sweep / k_chain_run
    int chain_threads = 256;                // CTA size of k_chain_run (256 x4, 512 x2 or 1024 x1 per SM)
    int cluster_p = 0;                      // > 1: few chains -> k_chain_run_cl, a cluster of cluster_p CTAs of CLUSTER_THREADS per chain
    size_t cluster_smem = 0;
    int n_sm = 148;
    std::vector<void*> allocs;
    std::vector<int> h_res_is_fw;          // per residue: 1 = inactive (framework)
    std::vector<int> h_kx, h_ky, h_kz;
    std::vector<int> h_row_kstart;         // first k of every (kx, ky) row (+ end)
    double* d_w = nullptr;                 // W(k) without the form factor (mmc_get_kvectors)
    double* d_out = nullptr;               // 16 doubles of results
    double* h_out = nullptr;               // pinned mirror
    void* h_stage = nullptr;               // pinned staging for state transfers
    size_t h_stage_bytes = 0;
    double* d_tape = nullptr; size_t tape_cap = 0;
    mmc_move_record* d_recs = nullptr; size_t recs_cap = 0;
    long long* h_tape_pos = nullptr;
    // cached CUDA graph of one batched step
    cudaGraphExec_t graph = nullptr; int graph_steps = 0;
    BatchCtl* d_ctl = nullptr;
    std::vector<int> h_count;              // host mirror of count[C][MMC_MAX_RES]
    int pending_kind = 0, pending_res = 0, pending_res2 = 0;
    int64_t launches = 0;
    unsigned long long steps_done = 0;     // Philox step counter across batches
    int pending_chain = -1;
    // per-block snapshots (mmc_snapshot_begin / _wait): device blob + pinned mirror per slot, copied out on copy_stream
    struct Snap { char* dev = nullptr; char* host = nullptr; cudaEvent_t taken = nullptr, landed = nullptr; bool pending = false; };
    Snap snap[2];
    size_t snap_off[6] = {0, 0, 0, 0, 0, 0}, snap_bytes = 0;   // count, com, offsets, energies, counters, error flag
    std::vector<int> snap_count[2];        // compacted [C][nres] counts handed to the caller
    cudaStream_t copy_stream = nullptr;
    // mmc_system_energy_all: the framework's share of the start-up, computed once per handle
    double2* d_sfw = nullptr; double* d_fwc = nullptr; bool fw_ready = false;
    double *d_pair_part = nullptr, *d_recip_part = nullptr, *d_eall = nullptr;
    // fused strict-mode trial (k_strict_trial): ticket counter, result block in host-mapped pinned memory, sequence number
    unsigned int* d_ticket = nullptr; StrictOut* h_sout = nullptr; StrictOut* d_sout = nullptr; unsigned long long strict_seq = 0;
};

template <typename T>
static cudaError_t dalloc(mmc_handle* h, T** p, size_t n) {
    void* q = nullptr;
    cudaError_t e = cudaMalloc(&q, (n ? n : 1) * sizeof(T));
    // zeroed on the NULL stream and waited for: the set-up copies that follow use both the NULL stream (cudaMemcpy) and the
    // library's own NON-BLOCKING stream, which the NULL stream does not order against
    if (e == cudaSuccess) { h->allocs.push_back(q); e = cudaMemset(q, 0, (n ? n : 1) * sizeof(T)); }
    if (e == cudaSuccess) e = cudaStreamSynchronize(0);   // set-up only: the zero fill is complete before any stream touches q
    *p = (T*)q;
    return e;
}

// The device raises MMC_ERRF_* bits in cs.error_flag and leaves them set; every host entry point that hands chain data
// back (mmc_synchronize, mmc_run_batch, mmc_download_chains, mmc_snapshot_wait) reports them through this function, so an
// overflow can never pass silently (the reference aborts there: CheckMoleculeIndex -> AbortRun, output_utils.f90:604).
// `clear`: the flag was read from the live device word (not from a snapshot blob) and is reset for the next report.
static int report_error_flag(mmc_handle* h, int flag, bool clear) {
    if (!flag) return 0;
    if (clear) cudaMemsetAsync(h->cs.error_flag, 0, sizeof(int), h->stream);
    if (flag & MMC_ERRF_TAPE) return fail("a chain consumed more uniforms than its tape holds (tape_stride too short for n_steps); the extra draws were 0.5");
    return fail("a creation move exceeded the slot capacity (molecule index exceeds the maximum allowed, NB_MAX_MOLECULE analogue); that move was skipped");
}

extern "C" const char* mmc_last_error(void) { return g_err.c_str(); }

// ComputeEwaldParameters / ComputeFourierIndices, prepare_utils.f90:169-214
static bool valid_k(int kx, int ky, int kz, const int kmax[3]) {
    const double a = (double)kx / (double)kmax[0], b = (double)ky / (double)kmax[1], c = (double)kz / (double)kmax[2];
    const double k2 = a * a + b * b + c * c;
    return std::fabs(k2) >= MMC_ERR && k2 <= 1.0;  // CheckValidReciprocalVector ewald_kvectors.f90:199
}
extern "C" int mmc_ewald_setup(double cutoff, double tolerance, const double box_len[3], double* alpha, int kmax[3], int* nk) {
    if (cutoff <= 0 || !box_len || !alpha || !kmax || !nk) return fail("mmc_ewald_setup: bad arguments");
    const double tol = std::fmin(std::fabs(tolerance), 0.5);
    const double s = std::sqrt(std::fabs(std::log(tol * cutoff)));
    const double a = std::sqrt(std::fabs(std::log(tol * cutoff * s))) / cutoff;
    const double p = std::sqrt(-std::log(tol * cutoff * (2.0 * s * a) * (2.0 * s * a)));
    for (int d = 0; d < 3; d++) kmax[d] = (int)std::lround(0.25 + box_len[d] * a * p / MMC_PI);
    int n = 0;
    for (int kx = 0; kx <= kmax[0]; kx++)
        for (int ky = -kmax[1]; ky <= kmax[1]; ky++)
            for (int kz = -kmax[2]; kz <= kmax[2]; kz++)
                if (!(kx == 0 && ky == 0 && kz == 0) && valid_k(kx, ky, kz, kmax)) n++;
    *alpha = a; *nk = n;
    return 0;
}

extern "C" void mmc_destroy(mmc_t* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->graph) cudaGraphExecDestroy(h->graph);
    for (void* p : h->allocs) cudaFree(p);
    if (h->d_tape) cudaFree(h->d_tape);
    if (h->d_recs) cudaFree(h->d_recs);
    if (h->h_out) cudaFreeHost(h->h_out);
    if (h->h_stage) cudaFreeHost(h->h_stage);
    if (h->h_tape_pos) cudaFreeHost(h->h_tape_pos);
    if (h->h_sout) cudaFreeHost(h->h_sout);
    for (auto& s : h->snap) {
        if (s.dev) cudaFree(s.dev);
        if (s.host) cudaFreeHost(s.host);
        if (s.taken) cudaEventDestroy(s.taken);
        if (s.landed) cudaEventDestroy(s.landed);
    }
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

static int ensure_stage(mmc_handle* h, size_t bytes) {
    if (bytes <= h->h_stage_bytes) return 0;
    if (h->h_stage) cudaFreeHost(h->h_stage);
    h->h_stage = nullptr; h->h_stage_bytes = 0;
    CK(cudaMallocHost(&h->h_stage, bytes));
    h->h_stage_bytes = bytes;
    return 0;
}

// uploads guests of one chain from packed (residue, molecule[, atom]) arrays that also contain the
// inactive residues (their entries are skipped)
static int upload_chain(mmc_handle* h, int c, const int* count, const double* com, const double* offsets) {
    const DevConst& dc = h->dc;
    const size_t nb = ((size_t)dc.g_mols * 3 + (size_t)dc.g_atoms * 6) * sizeof(double) + MMC_MAX_RES * sizeof(int);
    if (ensure_stage(h, nb)) return 1;
    CK(cudaStreamSynchronize(h->stream));
    double* s_com = (double*)h->h_stage; double* s_off = s_com + (size_t)dc.g_mols * 3;
    double* s_x = s_off + (size_t)dc.g_atoms * 3; double* s_y = s_x + dc.g_atoms; double* s_z = s_y + dc.g_atoms;
    int* s_cnt = (int*)(s_z + dc.g_atoms);
    memset(h->h_stage, 0, nb);
    size_t mo = 0, ao = 0;
    for (int r = 0; r < dc.nres; r++) {
        const int na = dc.natoms[r], n = count[r];
        if (dc.is_active[r]) {
            if (n > dc.cap[r]) return fail("residue %d: count %d exceeds capacity %d", r, n, dc.cap[r]);
            s_cnt[r] = n;
            for (int m = 0; m < n; m++) {
                const double* cm = com + (mo + m) * 3;
                for (int d = 0; d < 3; d++) s_com[((size_t)dc.g_mol_base[r] + m) * 3 + d] = cm[d];
                for (int a = 0; a < na; a++) {
                    const double* of = offsets + (ao + (size_t)m * na + a) * 3;
                    const size_t ia = dc.g_atom_base[r] + (size_t)m * na + a;
                    for (int d = 0; d < 3; d++) s_off[ia * 3 + d] = of[d];
                    s_x[ia] = cm[0] + of[0]; s_y[ia] = cm[1] + of[1]; s_z[ia] = cm[2] + of[2];
                }
            }
        }
        mo += n; ao += (size_t)n * na;
    }
    cudaStream_t st = h->stream;
    CK(cudaMemcpyAsync(h->cs.gcom + (size_t)c * dc.g_mols * 3, s_com, (size_t)dc.g_mols * 3 * sizeof(double), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(h->cs.goff + (size_t)c * dc.g_atoms * 3, s_off, (size_t)dc.g_atoms * 3 * sizeof(double), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(h->cs.gx + (size_t)c * dc.g_atoms, s_x, (size_t)dc.g_atoms * sizeof(double), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(h->cs.gy + (size_t)c * dc.g_atoms, s_y, (size_t)dc.g_atoms * sizeof(double), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(h->cs.gz + (size_t)c * dc.g_atoms, s_z, (size_t)dc.g_atoms * sizeof(double), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(h->cs.count + (size_t)c * MMC_MAX_RES, s_cnt, MMC_MAX_RES * sizeof(int), cudaMemcpyHostToDevice, st));
    CK(cudaMemsetAsync(h->cs.fw_pair + (size_t)c * dc.g_mols, 0xFF, (size_t)dc.g_mols * sizeof(double2), st));   // new geometry: cached framework sums are void
    CK(cudaStreamSynchronize(st));
    if (h->h_count.size() != (size_t)dc.n_chains * MMC_MAX_RES) h->h_count.assign((size_t)dc.n_chains * MMC_MAX_RES, 0);
    for (int r = 0; r < MMC_MAX_RES; r++) h->h_count[(size_t)c * MMC_MAX_RES + r] = s_cnt[r];
    return 0;
}

// Layout of the per-move sweeps (kernels_sweep.cuh): framework atoms grouped by a k-d split into compact
// tiles of 32 with a bounding box each, the packed k list, the Coulomb kernel table, the real-space reach, and
// the launch configuration of k_sweep / k_chain_run.
// Coulomb kernel table (common.cuh coul_kernel_tab): g(s) = erfc(alpha sqrt(s))/sqrt(s) on segments
// [2^e (1 + j/8), 2^e (1 + (j+1)/8)), degree-9 interpolation at Chebyshev nodes solved in long double.
// Domain: s_min (0.25 A^2, raised if more than MMC_GTAB_MAXSEG segments would be needed) to (xmax/alpha)^2.
// Degree-D interpolant of g(s) = erfc(alpha sqrt(s)) / sqrt(s) on segment j of octave e at Chebyshev nodes, solved in long double;
// returns the largest sampled |p - g| (long double evaluation of both).  coef[0..D].
static long double fit_coulomb_segment(double alpha, int e, int j, int D, long double* coef) {
    const int NS = 1 << MMC_GTAB_MBITS;
    const long double pi = 3.14159265358979323846264338327950288L;
    const long double s0 = ldexpl(1.0L + (long double)j / NS, e), w = ldexpl(1.0L / NS, e);
    auto g = [&](long double t) { const long double sv = s0 + (t + 1.0L) * 0.5L * w, rr = sqrtl(sv); return erfcl((long double)alpha * rr) / rr; };
    long double V[MMC_GTAB_DEG + 1][MMC_GTAB_DEG + 2];
    for (int r = 0; r <= D; r++) {
        const long double node = cosl(pi * (2 * r + 1) / (2.0L * (D + 1)));
        long double p = 1.0L;
        for (int c = 0; c <= D; c++) { V[r][c] = p; p *= node; }
        V[r][D + 1] = g(node);
    }
    for (int c = 0; c <= D; c++) {  // Gauss-Jordan with partial pivoting
        int piv = c;
        for (int r = c + 1; r <= D; r++) if (fabsl(V[r][c]) > fabsl(V[piv][c])) piv = r;
        for (int k = 0; k <= D + 1; k++) std::swap(V[c][k], V[piv][k]);
        for (int r = 0; r <= D; r++) if (r != c) {
            const long double f = V[r][c] / V[c][c];
            for (int k = c; k <= D + 1; k++) V[r][k] -= f * V[c][k];
        }
    }
    for (int c = 0; c <= D; c++) coef[c] = V[c][D + 1] / V[c][c];
    long double err = 0.0L;
    for (int i = 0; i <= 64; i++) {
        const long double t = -1.0L + i / 32.0L;
        long double p = coef[D];
        for (int c = D - 1; c >= 0; c--) p = p * t + coef[c];
        err = std::max(err, fabsl(p - g(t)));
    }
    return err;
}

// Coulomb kernel table (common.cuh coul_kernel_tab): g(s) on segments [2^e (1 + j/NS), 2^e (1 + (j+1)/NS)), NS = 2^MMC_GTAB_MBITS,
// degree <= MMC_GTAB_DEG.  Domain: s_min = 1 A^2 (raised if more than MMC_GTAB_MAXSEG segments would be needed) to (xmax/alpha)^2.
// The sums need ABSOLUTE accuracy (the near pairs' g ~ 0.1 .. 1 rounds at 1e-17 .. 1e-16), and g falls by 17 decades over the
// domain: far octaves reach that with fewer terms.  deg_start[0] / [1] = first segment from which degree D - 1 / D - 2 keeps the
// sampled interpolation error below MMC_GTAB_ABS_TOL in EVERY later segment (round 2: for the bench system's alpha the last
// octave, r = 16 .. 22.6 A with 65 % of the pairs in reach, needs degree 6: two coefficient loads fewer per pair).
#define MMC_GTAB_ABS_TOL 1.0e-17L
static void build_coulomb_table(double alpha, double xmax, std::vector<double>& tab, int* base, double* smin_out, double* smax_out, int deg_start[2]) {
    const int D = MMC_GTAB_DEG, M = MMC_GTAB_MBITS, NS = 1 << M;
    const double smax = (xmax / alpha) * (xmax / alpha);
    int e_hi = 0; std::frexp(smax, &e_hi);   // smax in [2^(e_hi-1), 2^e_hi)
    int e_lo = 0;   // s >= 1 A^2: closer pairs (overlapping insertions only) take coul_kernel_slow
    while ((e_hi - e_lo) * NS > MMC_GTAB_MAXSEG) e_lo++;
    const int nseg = (e_hi - e_lo) * NS;
    tab.assign((size_t)(D + 1) * MMC_GTAB_MAXSEG, 0.0);
    long double coef[MMC_GTAB_DEG + 1];
    // lowest admissible degree per segment, then made monotone from the far end
    std::vector<int> deg(nseg, D);
#ifndef MMC_GTAB_FIXED_DEGREE
    for (int seg = 0; seg < nseg; seg++)
        for (int dd = D - 2; dd < D; dd++)
            if (dd >= 1 && fit_coulomb_segment(alpha, e_lo + seg / NS, seg % NS, dd, coef) <= MMC_GTAB_ABS_TOL) { deg[seg] = dd; break; }
    for (int seg = nseg - 2; seg >= 0; seg--) deg[seg] = std::max(deg[seg], deg[seg + 1]);
#endif
    deg_start[0] = deg_start[1] = MMC_GTAB_MAXSEG;
    for (int seg = nseg - 1; seg >= 0; seg--) { if (deg[seg] <= D - 1) deg_start[0] = seg; if (deg[seg] <= D - 2) deg_start[1] = seg; }
    for (int seg = 0; seg < nseg; seg++) {
        const int dd = seg >= deg_start[1] ? D - 2 : seg >= deg_start[0] ? D - 1 : D;
        fit_coulomb_segment(alpha, e_lo + seg / NS, seg % NS, dd, coef);
        for (int c = 0; c <= dd; c++) tab[(size_t)c * MMC_GTAB_MAXSEG + seg] = (double)coef[c];
    }
    const double smin = std::ldexp(1.0, e_lo);
    union { double d; uint64_t u; } cv; cv.d = smin;
    *base = (int)(cv.u >> (52 - M));
    *smin_out = smin; *smax_out = std::ldexp(1.0, e_hi);
}

static int build_sweep_layout(mmc_handle* h, const mmc_desc* d) {
    DevConst& dc = h->dc;
    const size_t n = (size_t)dc.n_fw;
    std::vector<double> fx(n + 1), fy(n + 1), fz(n + 1), fq(n + 1); std::vector<int> ft(n + 1);
    if (n) {
        CK(cudaMemcpy(fx.data(), dc.fx, n * 8, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(fy.data(), dc.fy, n * 8, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(fz.data(), dc.fz, n * 8, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(fq.data(), dc.fq, n * 8, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(ft.data(), dc.ftype, n * 4, cudaMemcpyDeviceToHost));
    }
    std::vector<double> wx(n), wy(n), wz(n);  // wrapped into [lo, lo + L)
    std::vector<int> order(n);
    auto unit = [&](double u) { u -= std::floor(u); if (u >= 1.0) u = 0.0; return u; };
    for (size_t j = 0; j < n; j++) {
        // fractional coordinates of the (possibly tilted) cell, wrapped into [0, 1), and the lattice-equivalent point
        const double rz = fz[j] - dc.lo[2], ry = fy[j] - dc.lo[1], rx = fx[j] - dc.lo[0];
        const double fz_ = rz / dc.L[2], fy_ = (ry - dc.tilt[2] * fz_) / dc.L[1], fx_ = (rx - dc.tilt[0] * fy_ - dc.tilt[1] * fz_) / dc.L[0];
        const double ux = unit(fx_), uy = unit(fy_), uz = unit(fz_);
        wx[j] = dc.lo[0] + ux * dc.L[0] + uy * dc.tilt[0] + uz * dc.tilt[1]; wy[j] = dc.lo[1] + uy * dc.L[1] + uz * dc.tilt[2]; wz[j] = dc.lo[2] + uz * dc.L[2];
    }
    // Tiles = leaves of a k-d split of the wrapped positions: cut the longest axis of the current box so that the left part
    // holds a multiple of 32 atoms nearest to half.  Every tile but the last is FULL (one atom per lane) and compact: on the
    // 50 000-atom bench framework the half extents average 4.1 A (max 6.5) against 6.9 x 5.8 x 4.3 A (max 47: curve jumps) for
    // 32-atom runs of a Morton order, and a 24 A sphere meets 201 tiles instead of 253 (round 2; tools/tile_cull_stats.py).
    std::iota(order.begin(), order.end(), 0);
    {
        struct Range { size_t a, b; };
        std::vector<Range> stack; if (n) stack.push_back({0, n});
        const double* w[3] = {wx.data(), wy.data(), wz.data()};
        while (!stack.empty()) {
            const Range r = stack.back(); stack.pop_back();
            const size_t cnt = r.b - r.a;
            if (cnt <= TILE) continue;
            double lo3[3] = {1e300, 1e300, 1e300}, hi3[3] = {-1e300, -1e300, -1e300};
            for (size_t i = r.a; i < r.b; i++) for (int k = 0; k < 3; k++) { const double v = w[k][order[i]]; lo3[k] = std::min(lo3[k], v); hi3[k] = std::max(hi3[k], v); }
            int ax = 0; for (int k = 1; k < 3; k++) if (hi3[k] - lo3[k] > hi3[ax] - lo3[ax]) ax = k;
            size_t left = (size_t)std::max<long long>(1, std::llround((double)cnt / 2.0 / TILE)) * TILE;
            if (left >= cnt) left = cnt / 2;
            std::nth_element(order.begin() + r.a, order.begin() + r.a + left, order.begin() + r.b,
                             [&](int p, int q) { return w[ax][p] < w[ax][q] || (w[ax][p] == w[ax][q] && p < q); });
            stack.push_back({r.a, r.a + left}); stack.push_back({r.a + left, r.b});
        }
    }
    const int n_tiles = (int)((n + TILE - 1) / TILE);
    const size_t np = (size_t)n_tiles * TILE;
    std::vector<double> sx(np + 1, 0.0), sy(np + 1, 0.0), sz(np + 1, 0.0), sq(np + 1, 0.0), tbox((size_t)n_tiles * 6 + 1, 0.0);
    std::vector<int> st(np + 1, 0);
    for (int t = 0; t < n_tiles; t++) {
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        for (int l = 0; l < TILE; l++) {
            const size_t i = (size_t)t * TILE + l;
            if (i >= n) break;
            const int j = order[i];
            sx[i] = fx[j]; sy[i] = fy[j]; sz[i] = fz[j]; sq[i] = fq[j]; st[i] = ft[j];
            const double w[3] = {wx[j], wy[j], wz[j]};
            for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], w[k]); hi[k] = std::max(hi[k], w[k]); }
        }
        for (int k = 0; k < 3; k++) {
            tbox[(size_t)k * n_tiles + t] = 0.5 * (lo[k] + hi[k]);
            tbox[(size_t)(3 + k) * n_tiles + t] = 0.5 * (hi[k] - lo[k]) * (1.0 + 1e-12) + 1e-9;   // never under-estimate the box
        }
    }
    double *px, *py, *pz, *pq, *pb, *pe, *prq; int *pt, *prt;
    CK(dalloc(h, &px, np)); CK(dalloc(h, &py, np)); CK(dalloc(h, &pz, np)); CK(dalloc(h, &pq, np)); CK(dalloc(h, &pt, np));
    CK(dalloc(h, &pb, (size_t)n_tiles * 6));
    if (np) {
        CK(cudaMemcpy(px, sx.data(), np * 8, cudaMemcpyHostToDevice)); CK(cudaMemcpy(py, sy.data(), np * 8, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(pz, sz.data(), np * 8, cudaMemcpyHostToDevice)); CK(cudaMemcpy(pq, sq.data(), np * 8, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(pt, st.data(), np * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(pb, tbox.data(), (size_t)n_tiles * 6 * 8, cudaMemcpyHostToDevice));
    }
    dc.sx = px; dc.sy = py; dc.sz = pz; dc.sq = pq; dc.stype = pt; dc.tbox = pb; dc.n_tiles = n_tiles;
    {
        std::vector<double> tab;
        build_coulomb_table(dc.alpha, 7.0, tab, &dc.gtab_base, &dc.gtab_smin, &dc.gtab_smax, dc.gtab_deg_start);
        CK(dalloc(h, &pe, tab.size())); CK(cudaMemcpy(pe, tab.data(), tab.size() * 8, cudaMemcpyHostToDevice));
        dc.gtab = pe;
    }
    {   // per-residue charge / type rows for the guest part of the sweep
        std::vector<double> rq(MMC_MAX_RES * MMC_MAX_MOL_ATOMS, 0.0); std::vector<int> rt(MMC_MAX_RES * MMC_MAX_MOL_ATOMS, 0);
        for (int r = 0; r < dc.nres; r++) if (dc.is_active[r])
            for (int a = 0; a < dc.natoms[r]; a++) { rq[r * MMC_MAX_MOL_ATOMS + a] = dc.q[r][a]; rt[r * MMC_MAX_MOL_ATOMS + a] = dc.type[r][a]; }
        CK(dalloc(h, &prq, rq.size())); CK(dalloc(h, &prt, rt.size()));
        CK(cudaMemcpy(prq, rq.data(), rq.size() * 8, cudaMemcpyHostToDevice)); CK(cudaMemcpy(prt, rt.data(), rt.size() * 4, cudaMemcpyHostToDevice));
        dc.res_q = prq; dc.res_type = prt;
    }
    {   // Reciprocal work items.  The k list (PrecomputeValidReciprocalVectors, ewald_kvectors.f90:44-103) is
        // ordered kx, ky, kz with kz fastest, so a (kx, ky) row holds kz = -m..m contiguously (row (0,0) lacks
        // kz = 0) and +kz / -kz share X(kx) Y(ky), |Z(kz)| and f W.  One item = {row, kz >= 0} -> k(+kz), k(-kz).
        if (dc.kmax[0] > 1023 || dc.kmax[1] > 1023 || dc.kmax[2] > 1023) return fail("mmc_init: kmax above 1023 is not supported");
        const int ny = 2 * dc.kmax[1] + 1, nrows = (dc.kmax[0] + 1) * ny;
        std::vector<int> rs(nrows + 1, dc.nk);
        for (int i = dc.nk - 1; i >= 0; i--) rs[h->h_kx[i] * ny + h->h_ky[i] + dc.kmax[1]] = i;
        for (int r = nrows - 1; r >= 0; r--) if (rs[r] > rs[r + 1]) rs[r] = rs[r + 1];   // empty rows point at the next one
        std::vector<int2> items; items.reserve(dc.nk / 2 + nrows);
        std::vector<int> irs(nrows + 1, 0);
        for (int r = 0; r < nrows; r++) {
            irs[r] = (int)items.size();
            const int k0 = rs[r], k1 = rs[r + 1];
            for (int kpl = k0; kpl < k1; kpl++) {
                const int kz = h->h_kz[kpl];
                if (kz < 0) continue;
                uint32_t flag = 0;
                if (kz > 0) {
                    int kmi = kpl - 2 * kz;
                    if (kmi < k0 || h->h_kz[kmi] != -kz) { kmi++; flag = 1u; }
                    if (kmi < k0 || kmi >= k1 || h->h_kz[kmi] != -kz || h->h_kx[kmi] != h->h_kx[kpl] || h->h_ky[kmi] != h->h_ky[kpl])
                        return fail("mmc_init: k list is not symmetric in kz inside row %d", r);
                }
                const uint32_t pack = (uint32_t)h->h_kx[kpl] | ((uint32_t)(h->h_ky[kpl] + dc.kmax[1]) << 10) | ((uint32_t)kz << 21) | (flag << 31);
                items.push_back(make_int2((int)pack, kpl));
            }
        }
        irs[nrows] = (int)items.size();
        size_t covered = 0;
        for (const int2& it : items) covered += (((uint32_t)it.x >> 21) & 1023u) ? 2 : 1;
        if (covered != (size_t)dc.nk) return fail("mmc_init: reciprocal items cover %zu of %d k-vectors", covered, dc.nk);
        int2* pit; int *pirs, *prs;
        CK(dalloc(h, &prs, rs.size())); CK(cudaMemcpy(prs, rs.data(), rs.size() * 4, cudaMemcpyHostToDevice));
        dc.row_kstart = prs; h->h_row_kstart = rs;
        CK(dalloc(h, &pit, items.size() + 1)); CK(cudaMemcpy(pit, items.data(), items.size() * sizeof(int2), cudaMemcpyHostToDevice));
        CK(dalloc(h, &pirs, irs.size())); CK(cudaMemcpy(pirs, irs.data(), irs.size() * 4, cudaMemcpyHostToDevice));
        dc.kitem = pit; dc.item_row_start = pirs; dc.n_items = (int)items.size();
    }
    // reach of the real-space sweep
    bool any_charge = false;
    for (size_t j = 0; j < n; j++) if (std::fabs(fq[j]) >= MMC_ERR) any_charge = true;
    for (int r = 0; r < dc.nres; r++) if (dc.is_active[r]) for (int a = 0; a < dc.natoms[r]; a++) if (std::fabs(dc.q[r][a]) >= MMC_ERR) any_charge = true;
    dc.rc2 = dc.rc * dc.rc; dc.alpha2 = dc.alpha * dc.alpha;
    double xcut = d->coulomb_xcut == 0.0 ? 6.0 : d->coulomb_xcut;
    if (xcut > 7.0) xcut = 7.0;
    if (xcut < 0.0) { dc.reach = 1e200; dc.reach2 = INFINITY; }
    else { dc.reach = any_charge ? std::max(dc.rc, xcut / dc.alpha) : dc.rc; dc.reach2 = dc.reach * dc.reach; }
    // launch configuration
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, h->device));
    h->n_sm = prop.multiProcessorCount;
    const int work = std::max((dc.nk + 511) / 512, (n_tiles + dc.g_atoms / TILE + 15) / 16);
    dc.sweep_parts = std::max(1, std::min(2 * h->n_sm, work));
    size_t off[12];
    h->sweep_smem = sweep_smem_layout(dc.max_nat, dc.ntypes, dc.kmax, SWEEP_THREADS, off);
    h->chain_threads = dc.n_chains >= 3 * h->n_sm ? 256 : dc.n_chains >= 3 * h->n_sm / 2 ? 512 : 1024;
    h->chain_smem = sweep_smem_layout(dc.max_nat, dc.ntypes, dc.kmax, h->chain_threads, off);
    if (h->chain_smem > 200 * 1024) return fail("kmax=(%d,%d,%d) with %d-atom molecules needs %zu B of shared memory for the phase tables",
                                                 dc.kmax[0], dc.kmax[1], dc.kmax[2], dc.max_nat, h->chain_smem);
    // The opt-in limit is a property of the FUNCTION, shared by every handle of the process: set it to the ceiling once instead
    // of to this handle's need (a second, smaller system must not lower it under a live handle's launches -- r2e: "invalid
    // argument" on the bench system's strict-mode launch after a ZIF-8 handle had been created)
    const int smem_cap = 200 * 1024;
    if (h->sweep_smem > (size_t)smem_cap) return fail("kmax=(%d,%d,%d) with %d-atom molecules needs %zu B of shared memory for the phase tables",
                                                       dc.kmax[0], dc.kmax[1], dc.kmax[2], dc.max_nat, h->sweep_smem);
    CK(cudaFuncSetAttribute(k_sweep, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
    CK(cudaFuncSetAttribute(k_strict_trial, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
    CK(cudaFuncSetAttribute(k_chain_run<256, 4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
    CK(cudaFuncSetAttribute(k_chain_run<512, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
    CK(cudaFuncSetAttribute(k_chain_run<1024, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
    CK(cudaFuncSetAttribute(k_chain_run<256, 4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
    CK(cudaFuncSetAttribute(k_chain_run<512, 2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
    CK(cudaFuncSetAttribute(k_chain_run<1024, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
    {   // Few chains: a thread-block cluster per chain (k_chain_run_cl).  Measured on B200 (tools/few_chains.py,
        // profiles/r2n_few_chains.log): it wins only while chains <= SMs / 2 (64 chains: +9 % at P = 4; 32 chains: +26 % at P = 8;
        // c4_large_k at 64 chains: +31 %), with P = the largest of 8, 4, 2 that keeps P x chains CTAs resident (two 512-thread
        // CTAs per SM); at 128 chains and above one CTA per chain is faster.  MANIAC_CLUSTER=<P> overrides (0 = off).
        int P = 0;
        if (2 * dc.n_chains <= h->n_sm) for (int q = 8; q >= 2; q >>= 1) if ((long long)q * dc.n_chains <= 2LL * h->n_sm) { P = q; break; }
        if (const char* e = getenv("MANIAC_CLUSTER")) { P = atoi(e); if (P != 2 && P != 4 && P != 8) P = 0; }
        h->cluster_p = P > 1 ? P : 0;
        h->cluster_smem = sweep_smem_layout(dc.max_nat, dc.ntypes, dc.kmax, CLUSTER_THREADS, off);
        CK(cudaFuncSetAttribute(k_chain_run_cl<CLUSTER_THREADS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
        CK(cudaFuncSetAttribute(k_chain_run_cl<CLUSTER_THREADS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_cap));
    }
    CK(dalloc(h, &h->cs.part, (size_t)dc.n_chains * dc.sweep_parts * 6));
    return 0;
}

extern "C" int mmc_init(mmc_t** out, const mmc_desc* d) {
    if (!out || !d) return fail("mmc_init: null argument");
    *out = nullptr;
    if (d->nres < 1 || d->nres > MMC_MAX_RES) return fail("mmc_init: nres=%d outside [1,%d]", d->nres, MMC_MAX_RES);
    if (d->n_chains < 1) return fail("mmc_init: n_chains must be >= 1");
    if (d->ntypes < 1) return fail("mmc_init: ntypes must be >= 1");
    if (!d->natoms || !d->is_active || !d->count || !d->charges || !d->types || !d->sigma || !d->epsilon) return fail("mmc_init: null array in mmc_desc");
    for (int r = 0; r < d->nres; r++) {
        if (d->natoms[r] < 1) return fail("mmc_init: residue %d has %d atoms", r, d->natoms[r]);
        if (d->natoms[r] > d->max_atoms) return fail("mmc_init: residue %d has %d atoms but max_atoms (leading dimension of charges/types) is %d", r, d->natoms[r], d->max_atoms);
        if (d->count[r] < 0) return fail("mmc_init: residue %d has a negative molecule count", r);
    }
    for (int k = 0; k < 3; k++) {
        if (!(d->box_len[k] > 0.0)) return fail("mmc_init: box_len[%d] must be positive (cell edge lengths lx, ly, lz; tilt factors go in mmc_desc.tilt)", k);
        if (d->kmax[k] < 1) return fail("mmc_init: kmax[%d] must be >= 1", k);
    }
    if (!(d->alpha > 0.0) || !(d->cutoff > 0.0) || !(d->temperature > 0.0)) return fail("mmc_init: alpha, cutoff and temperature must be positive");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail("mmc_init: no CUDA device available (this library has no CPU fallback)");
    if (d->device < 0 || d->device >= ndev) return fail("mmc_init: device %d out of range (%d devices)", d->device, ndev);
    CK(cudaSetDevice(d->device));
    mmc_handle* h = new mmc_handle();
    h->device = d->device; h->init_sk = d->init_sk;
    if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) { delete h; return fail("cudaStreamCreate failed"); }
    DevConst& dc = h->dc;
    memset(&dc, 0, sizeof dc);
    dc.nres = d->nres; dc.ntypes = d->ntypes; dc.n_chains = d->n_chains;
    dc.deletion_mode = d->deletion_mode; dc.use_table = d->use_table;
    dc.fw_cache = 1;   // MANIAC_FW_CACHE=0: re-sum the framework part of the old geometry in every trial, as the reference does (A/B switch)
    if (const char* e = getenv("MANIAC_FW_CACHE")) dc.fw_cache = atoi(e) != 0;
    dc.shard_count = d->shard_count > 1 ? d->shard_count : 1;
    dc.shard_rank = d->shard_count > 1 ? d->shard_rank : 0;
    if (dc.shard_rank < 0 || dc.shard_rank >= dc.shard_count) { delete h; return fail("mmc_init: shard_rank %d outside [0,%d)", d->shard_rank, dc.shard_count); }
    dc.alpha = d->alpha; dc.rc = d->cutoff;
    dc.coul_const = MMC_EPS0_INV_eVA / MMC_KB_eVK;
    dc.p_translation = d->p_translation; dc.p_rotation = d->p_rotation; dc.p_swap = d->p_swap;
    dc.volume = d->box_len[0] * d->box_len[1] * d->box_len[2];
    dc.recip_const = MMC_EPS0_INV_eVA / MMC_KB_eVK * MMC_TWOPI / dc.volume;
    int kmaxmax = 0;
    for (int k = 0; k < 3; k++) {
        dc.lo[k] = d->box_lo[k]; dc.L[k] = d->box_len[k]; dc.invL[k] = 1.0 / d->box_len[k]; dc.kmax[k] = d->kmax[k];
        if (d->kmax[k] > kmaxmax) kmaxmax = d->kmax[k];
    }
    dc.kstride = kmaxmax + 1;
    for (int k = 0; k < 3; k++) dc.tilt[k] = d->tilt[k];
    dc.tri = d->tilt[0] != 0.0 || d->tilt[1] != 0.0 || d->tilt[2] != 0.0;
    { const double m = std::fmin(dc.L[0], std::fmin(dc.L[1], dc.L[2])); dc.half_min2 = 0.25 * m * m; }
#define BAIL(...) do { fail(__VA_ARGS__); std::string keep_ = g_err; mmc_destroy(h); g_err = keep_; return 1; } while (0)
#define CKB(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) BAIL("%s failed: %s", #call, cudaGetErrorString(e_)); } while (0)
    // residues
    int g_atoms = 0, g_mols = 0; size_t n_fw = 0; int n_fw_mols = 0;
    for (int r = 0; r < d->nres; r++) {
        dc.natoms[r] = d->natoms[r]; dc.is_active[r] = d->is_active[r] == 1;
        if (dc.is_active[r]) {
            if (d->natoms[r] > MMC_MAX_MOL_ATOMS) BAIL("active residue %d has %d atoms (> MMC_MAX_MOL_ATOMS=%d)", r, d->natoms[r], MMC_MAX_MOL_ATOMS);
            int cap = d->capacity ? d->capacity[r] : d->count[r];
            if (cap < d->count[r]) cap = d->count[r];
            if (cap < 1) cap = 1;
            dc.cap[r] = cap;
            dc.active_list[dc.n_active++] = r;
            dc.g_atom_base[r] = g_atoms; dc.g_mol_base[r] = g_mols;
            g_atoms += cap * d->natoms[r]; g_mols += cap;
            double se = 0.0;  // ComputeEwaldSelfInteractionSingleMol ewald_energy.f90:308-336 (constant per type)
            for (int a = 0; a < d->natoms[r]; a++) {
                const double q = d->charges[(size_t)r * d->max_atoms + a];
                const int ty = d->types[(size_t)r * d->max_atoms + a];
                if (ty < 1 || ty > d->ntypes) BAIL("residue %d atom %d: type %d outside [1,%d]", r, a, ty, d->ntypes);
                dc.q[r][a] = q; dc.type[r][a] = ty - 1;
                if (std::fabs(q) < MMC_ERR) continue;
                se = se - d->alpha / std::sqrt(MMC_PI) * (q * q);
            }
            dc.self_e[r] = se * dc.coul_const;
        } else {
            dc.cap[r] = d->count[r];
            n_fw += (size_t)d->count[r] * d->natoms[r]; n_fw_mols += d->count[r];
        }
    }
    dc.g_atoms = g_atoms > 0 ? g_atoms : 1; dc.g_mols = g_mols > 0 ? g_mols : 1; dc.n_fw = (int)n_fw; h->n_fw_mols = n_fw_mols;
    // framework SoA (absolute positions com + offset, as ComputeDistance forms them, geometry_utils.f90:379-382)
    {
        std::vector<double> fx(n_fw + 1), fy(n_fw + 1), fz(n_fw + 1), fq(n_fw + 1); std::vector<int> ft(n_fw + 1), fm(n_fw + 1);
        size_t mo = 0, ao = 0, j = 0; int molid = 0;
        for (int r = 0; r < d->nres; r++) {
            const int na = d->natoms[r], n = d->count[r];
            if (!dc.is_active[r])
                for (int m = 0; m < n; m++, molid++)
                    for (int a = 0; a < na; a++, j++) {
                        const double* cm = d->com + (mo + m) * 3; const double* of = d->offsets + (ao + (size_t)m * na + a) * 3;
                        fx[j] = cm[0] + of[0]; fy[j] = cm[1] + of[1]; fz[j] = cm[2] + of[2];
                        fq[j] = d->charges[(size_t)r * d->max_atoms + a];
                        const int ty = d->types[(size_t)r * d->max_atoms + a];
                        if (ty < 1 || ty > d->ntypes) BAIL("residue %d atom %d: type %d outside [1,%d]", r, a, ty, d->ntypes);
                        ft[j] = ty - 1; fm[j] = molid;
                    }
            mo += n; ao += (size_t)n * na;
        }
        double *px, *py, *pz, *pq; int *pt, *pm;
        CKB(dalloc(h, &px, n_fw)); CKB(dalloc(h, &py, n_fw)); CKB(dalloc(h, &pz, n_fw)); CKB(dalloc(h, &pq, n_fw));
        CKB(dalloc(h, &pt, n_fw)); CKB(dalloc(h, &pm, n_fw));
        if (n_fw) {
            CKB(cudaMemcpy(px, fx.data(), n_fw * 8, cudaMemcpyHostToDevice)); CKB(cudaMemcpy(py, fy.data(), n_fw * 8, cudaMemcpyHostToDevice));
            CKB(cudaMemcpy(pz, fz.data(), n_fw * 8, cudaMemcpyHostToDevice)); CKB(cudaMemcpy(pq, fq.data(), n_fw * 8, cudaMemcpyHostToDevice));
            CKB(cudaMemcpy(pt, ft.data(), n_fw * 4, cudaMemcpyHostToDevice)); CKB(cudaMemcpy(pm, fm.data(), n_fw * 4, cudaMemcpyHostToDevice));
        }
        dc.fx = px; dc.fy = py; dc.fz = pz; dc.fq = pq; dc.ftype = pt; dc.fmol = pm;
    }
    // LJ table by atom type: {sigma, 4*epsilon}
    {
        const size_t nt2 = (size_t)d->ntypes * d->ntypes;
        std::vector<double2> lj(nt2);
        for (size_t i = 0; i < nt2; i++) lj[i] = make_double2(d->sigma[i], 4.0 * d->epsilon[i]);
        double2* p; CKB(dalloc(h, &p, nt2)); CKB(cudaMemcpy(p, lj.data(), nt2 * sizeof(double2), cudaMemcpyHostToDevice));
        dc.lj = p;
    }
    // k-vector list, PrecomputeValidReciprocalVectors ewald_kvectors.f90:44-99 (kx outer, kz inner)
    {
        std::vector<int4> kv;
        for (int kx = 0; kx <= dc.kmax[0]; kx++)
            for (int ky = -dc.kmax[1]; ky <= dc.kmax[1]; ky++)
                for (int kz = -dc.kmax[2]; kz <= dc.kmax[2]; kz++) {
                    if (kx == 0 && ky == 0 && kz == 0) continue;
                    if (!valid_k(kx, ky, kz, dc.kmax)) continue;
                    kv.push_back(make_int4(kx, ky, kz, 0)); h->h_kx.push_back(kx); h->h_ky.push_back(ky); h->h_kz.push_back(kz);
                }
        dc.nk = (int)kv.size();
        if (dc.nk == 0) BAIL("mmc_init: empty k-vector set");
        dc.n_k_tiles = (dc.nk + RECIP_THREADS - 1) / RECIP_THREADS;
        int4* p; CKB(dalloc(h, &p, kv.size())); CKB(cudaMemcpy(p, kv.data(), kv.size() * sizeof(int4), cudaMemcpyHostToDevice));
        dc.kvec = p;
        double *fw, *w; CKB(dalloc(h, &fw, dc.nk)); CKB(dalloc(h, &w, dc.nk));
        dc.fw = fw;
        k_init_weights<<<(dc.nk + 255) / 256, 256, 0, h->stream>>>(dc, fw, w); h->launches++;
        h->d_w = w;
    }
    // tables (PrecomputeTable tabulated_utils.f90:10): built every run like the reference does
    {
        double *t0, *t1, *t2; CKB(dalloc(h, &t0, 5001)); CKB(dalloc(h, &t1, 5001)); CKB(dalloc(h, &t2, 5001));
        k_init_tables<<<(5001 + 255) / 256, 256, 0, h->stream>>>(dc.alpha, dc.rc, t0, t1, t2); h->launches++;
        dc.tab_erfc = t0; dc.tab_r6 = t1; dc.tab_r12 = t2; dc.tab_dx = dc.rc / 5000.0;
    }
    // per-chain state
    ChainState& cs = h->cs;
    const size_t C = dc.n_chains;
    CKB(dalloc(h, &cs.count, C * MMC_MAX_RES));
    CKB(dalloc(h, &cs.gx, C * dc.g_atoms)); CKB(dalloc(h, &cs.gy, C * dc.g_atoms)); CKB(dalloc(h, &cs.gz, C * dc.g_atoms));
    CKB(dalloc(h, &cs.gcom, C * dc.g_mols * 3)); CKB(dalloc(h, &cs.goff, C * dc.g_atoms * 3));
    CKB(dalloc(h, &cs.A, C * 2 * dc.nk)); CKB(dalloc(h, &cs.cur, C));
    CKB(dalloc(h, &cs.energy, C * MMC_NE)); CKB(dalloc(h, &cs.fug, C * MMC_MAX_RES));
    CKB(dalloc(h, &cs.temp, C)); CKB(dalloc(h, &cs.tstep, C)); CKB(dalloc(h, &cs.rstep, C));
    CKB(dalloc(h, &cs.counters, C)); CKB(dalloc(h, &cs.trial, C));
    CKB(dalloc(h, &cs.fw_pair, C * dc.g_mols)); CKB(cudaMemset(cs.fw_pair, 0xFF, C * dc.g_mols * sizeof(double2)));   // all NaN: nothing known yet
    {   // reservoir pools (mmc_desc.reservoir_*): one per chain, every chain starts from the same pool
        dc.has_reservoir = (d->reservoir_count && d->reservoir_offsets) ? 1 : 0;
        int ra = 0;
        for (int r = 0; r < dc.nres; r++) {
            dc.r_atom_base[r] = ra; dc.r_cap[r] = 0;
            if (dc.has_reservoir && dc.is_active[r]) { dc.r_cap[r] = d->reservoir_count[r] + dc.cap[r] + 1; ra += dc.r_cap[r] * dc.natoms[r]; }
        }
        dc.r_atoms = ra;
        CKB(dalloc(h, &cs.rcount, C * MMC_MAX_RES)); CKB(dalloc(h, &cs.roff, C * (size_t)std::max(ra, 1) * 3));
        if (dc.has_reservoir) {
            std::vector<int> rc(C * MMC_MAX_RES, 0); std::vector<double> ro(C * (size_t)std::max(ra, 1) * 3, 0.0);
            size_t src = 0;
            for (int r = 0; r < dc.nres; r++) {
                const size_t n = (size_t)d->reservoir_count[r] * dc.natoms[r] * 3;
                if (dc.is_active[r]) for (size_t c = 0; c < C; c++) {
                    rc[c * MMC_MAX_RES + r] = d->reservoir_count[r];
                    memcpy(ro.data() + (c * (size_t)ra + dc.r_atom_base[r]) * 3, d->reservoir_offsets + src, n * sizeof(double));
                }
                src += n;
            }
            CKB(cudaMemcpy(cs.rcount, rc.data(), rc.size() * sizeof(int), cudaMemcpyHostToDevice));
            CKB(cudaMemcpy(cs.roff, ro.data(), ro.size() * sizeof(double), cudaMemcpyHostToDevice));
        }
    }
    CKB(dalloc(h, &cs.tape_pos, C)); CKB(dalloc(h, &cs.chain_id, C)); CKB(dalloc(h, &cs.error_flag, 1)); CKB(dalloc(h, &cs.pair_count, 10));
    CKB(dalloc(h, &h->d_out, 16)); CKB(dalloc(h, &h->d_ctl, 1));
    CKB(cudaMallocHost((void**)&h->h_out, 16 * sizeof(double)));
    CKB(cudaMallocHost((void**)&h->h_tape_pos, C * sizeof(long long)));
    CKB(dalloc(h, &h->d_ticket, 1));
    CKB(cudaHostAlloc((void**)&h->h_sout, sizeof(StrictOut), cudaHostAllocMapped));
    memset(h->h_sout, 0, sizeof(StrictOut));
    CKB(cudaHostGetDevicePointer((void**)&h->d_sout, h->h_sout, 0));
    memset(h->h_tape_pos, 0, C * sizeof(long long));
    {
        std::vector<unsigned int> ids(C);
        for (size_t c = 0; c < C; c++) ids[c] = (unsigned int)c;
        CKB(cudaMemcpy(cs.chain_id, ids.data(), C * sizeof(unsigned int), cudaMemcpyHostToDevice));
    }
    {
        std::vector<double> fug(C * MMC_MAX_RES, 0.0), tmp(C, d->temperature), ts(C, d->translation_step), rs(C, d->rotation_step);
        for (size_t c = 0; c < C; c++) for (int r = 0; r < d->nres; r++) fug[c * MMC_MAX_RES + r] = d->fugacity ? d->fugacity[r] : 0.0;
        CKB(cudaMemcpy(cs.fug, fug.data(), fug.size() * 8, cudaMemcpyHostToDevice));
        CKB(cudaMemcpy(cs.temp, tmp.data(), C * 8, cudaMemcpyHostToDevice));
        CKB(cudaMemcpy(cs.tstep, ts.data(), C * 8, cudaMemcpyHostToDevice));
        CKB(cudaMemcpy(cs.rstep, rs.data(), C * 8, cudaMemcpyHostToDevice));
    }
    for (size_t c = 0; c < C; c++)
        if (upload_chain(h, (int)c, d->count, d->com, d->offsets)) { std::string keep = g_err; mmc_destroy(h); g_err = keep; return 1; }
    // scratch of the single-point energy
    SysScratch& ss = h->ss;
    memset(&ss, 0, sizeof ss);
    ss.n = dc.n_fw + dc.g_atoms; ss.n_tiles = (ss.n + SYS_THREADS - 1) / SYS_THREADS;
    CKB(dalloc(h, &ss.x, ss.n)); CKB(dalloc(h, &ss.y, ss.n)); CKB(dalloc(h, &ss.z, ss.n)); CKB(dalloc(h, &ss.q, ss.n));
    CKB(dalloc(h, &ss.type, ss.n)); CKB(dalloc(h, &ss.mol, ss.n));
    CKB(dalloc(h, &ss.tile_min, ss.n_tiles)); CKB(dalloc(h, &ss.tile_max, ss.n_tiles));
    const size_t npart = std::max((size_t)ss.n_tiles * ss.n_tiles, (size_t)dc.n_k_tiles) * 2;
    CKB(dalloc(h, &ss.part, npart));
    h->sk_chunk = SK_CHUNK;
    while (h->sk_chunk > 1 && (size_t)h->sk_chunk * 3 * dc.kstride * sizeof(double2) > 48 * 1024) h->sk_chunk /= 2;
    h->sk_smem = (size_t)h->sk_chunk * 3 * dc.kstride * sizeof(double2);
    const int n_chunks = (ss.n + h->sk_chunk - 1) / h->sk_chunk;
    ss.n_super = n_chunks < 128 ? n_chunks : 128;
    CKB(dalloc(h, &ss.sk_part, (size_t)ss.n_super * dc.nk));
    CKB(dalloc(h, &ss.result, 8));
    int max_nat = 1;
    for (int i = 0; i < dc.n_active; i++) if (dc.natoms[dc.active_list[i]] > max_nat) max_nat = dc.natoms[dc.active_list[i]];
    dc.max_nat = max_nat;
    if (dc.kmax[0] > 1023 || dc.kmax[1] > 511 || dc.kmax[2] > 511) BAIL("kmax too large for the packed k-vector list");
    if (build_sweep_layout(h, d)) { std::string keep = g_err; mmc_destroy(h); g_err = keep; return 1; }
    CKB(cudaDeviceSynchronize());   // set-up used both the NULL stream (copies) and the library stream (kernels)
    CKB(cudaGetLastError());
    *out = h;
    return 0;
}

// ---------------------------------------------------------------------------------------------
extern "C" int mmc_num_kvectors(const mmc_t* h) { return h ? h->dc.nk : -1; }

extern "C" int mmc_get_kvectors(mmc_t* h, int* kx, int* ky, int* kz, double* weight, double* form) {
    if (!h) return fail("null handle");
    CK(cudaSetDevice(h->device));
    const int n = h->dc.nk;
    for (int i = 0; i < n; i++) {
        if (kx) kx[i] = h->h_kx[i];
        if (ky) ky[i] = h->h_ky[i];
        if (kz) kz[i] = h->h_kz[i];
        if (form) form[i] = h->h_kx[i] == 0 ? 1.0 : 2.0;  // ComputeSymmetryFormFactor ewald_kvectors.f90:167
    }
    if (weight) { CK(cudaStreamSynchronize(h->stream)); CK(cudaMemcpy(weight, h->d_w, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost)); }
    return 0;
}

extern "C" int mmc_set_chain_params(mmc_t* h, int c, const double* fugacity, double temperature) {
    CHECK_CHAIN(h, c);
    CK(cudaSetDevice(h->device));
    if (!(temperature > 0.0)) return fail("temperature must be positive");
    double f[MMC_MAX_RES] = {0};
    for (int r = 0; r < h->dc.nres; r++) f[r] = fugacity ? fugacity[r] : 0.0;
    // on the library's (non-blocking) stream, then waited for: ordered against the kernels before and after
    if (fugacity) CK(cudaMemcpyAsync(h->cs.fug + (size_t)c * MMC_MAX_RES, f, sizeof f, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->cs.temp + c, &temperature, sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int mmc_adjust_step_sizes(mmc_t* h) {
    if (!h) return fail("null handle");
    CK(cudaSetDevice(h->device));
    k_adjust_step_sizes<<<(h->dc.n_chains + 127) / 128, 128, 0, h->stream>>>(h->dc, h->cs);
    h->launches++;
    CK(cudaGetLastError());
    return 0;
}
extern "C" int mmc_get_step_sizes(mmc_t* h, int c, double* tstep, double* rstep) {
    CHECK_CHAIN(h, c);
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    if (tstep) CK(cudaMemcpy(tstep, h->cs.tstep + c, sizeof(double), cudaMemcpyDeviceToHost));
    if (rstep) CK(cudaMemcpy(rstep, h->cs.rstep + c, sizeof(double), cudaMemcpyDeviceToHost));
    return 0;
}
extern "C" int mmc_set_chain_id(mmc_t* h, int c, uint32_t id) {
    CHECK_CHAIN(h, c);
    CK(cudaSetDevice(h->device));
    CK(cudaMemcpyAsync(h->cs.chain_id + c, &id, sizeof id, cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}
extern "C" int mmc_set_step_sizes(mmc_t* h, int c, double tstep, double rstep) {
    CHECK_CHAIN(h, c);
    CK(cudaSetDevice(h->device));
    CK(cudaMemcpyAsync(h->cs.tstep + c, &tstep, sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->cs.rstep + c, &rstep, sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return 0;
}

extern "C" int mmc_set_chain_state(mmc_t* h, int c, const int* count, const double* com, const double* offsets) {
    CHECK_CHAIN(h, c);
    if (!count || !com || !offsets) return fail("mmc_set_chain_state: null argument");
    CK(cudaSetDevice(h->device));
    return upload_chain(h, c, count, com, offsets);
}

// downloads one chain; inactive residues are not written (the host already has them): their
// entries in the packed arrays are skipped over
static int download_chain(mmc_handle* h, int c, int* count, double* com, double* offsets) {
    const DevConst& dc = h->dc;
    const size_t nb = ((size_t)dc.g_mols * 3 + (size_t)dc.g_atoms * 3) * sizeof(double) + MMC_MAX_RES * sizeof(int);
    if (ensure_stage(h, nb)) return 1;
    double* s_com = (double*)h->h_stage; double* s_off = s_com + (size_t)dc.g_mols * 3; int* s_cnt = (int*)(s_off + (size_t)dc.g_atoms * 3);
    cudaStream_t st = h->stream;
    CK(cudaMemcpyAsync(s_cnt, h->cs.count + (size_t)c * MMC_MAX_RES, MMC_MAX_RES * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (com) CK(cudaMemcpyAsync(s_com, h->cs.gcom + (size_t)c * dc.g_mols * 3, (size_t)dc.g_mols * 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (offsets) CK(cudaMemcpyAsync(s_off, h->cs.goff + (size_t)c * dc.g_atoms * 3, (size_t)dc.g_atoms * 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    size_t mo = 0, ao = 0;
    for (int r = 0; r < dc.nres; r++) {
        const int na = dc.natoms[r];
        const int n = dc.is_active[r] ? s_cnt[r] : dc.cap[r];
        if (count) count[r] = n;
        if (dc.is_active[r]) {
            if (com) memcpy(com + mo * 3, s_com + (size_t)dc.g_mol_base[r] * 3, (size_t)n * 3 * sizeof(double));
            if (offsets) memcpy(offsets + ao * 3, s_off + (size_t)dc.g_atom_base[r] * 3, (size_t)n * na * 3 * sizeof(double));
        }
        mo += n; ao += (size_t)n * na;
    }
    return 0;
}

extern "C" int mmc_get_chain_state(mmc_t* h, int c, int* count, double* com, double* offsets) {
    CHECK_CHAIN(h, c);
    CK(cudaSetDevice(h->device));
    return download_chain(h, c, count, com, offsets);
}

extern "C" int mmc_sync_to_host(mmc_t* h, int c, int* count, double* com, double* offsets, double energies[MMC_NE]) {
    CHECK_CHAIN(h, c);
    CK(cudaSetDevice(h->device));
    if (download_chain(h, c, count, com, offsets)) return 1;
    if (energies) CK(cudaMemcpy(energies, h->cs.energy + (size_t)c * MMC_NE, MMC_NE * sizeof(double), cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int mmc_chain_capacity(const mmc_t* h, int* g_mols, int* g_atoms) {
    if (!h) return fail("null handle");
    if (g_mols) *g_mols = h->dc.g_mols;
    if (g_atoms) *g_atoms = h->dc.g_atoms;
    return 0;
}

// the host arrays hold nres counts per chain, the device MMC_MAX_RES: one strided 2-D copy
extern "C" int mmc_upload_chains(mmc_t* h, const int* count, const double* com, const double* offsets) {
    if (!h) return fail("null handle");
    if (!count || !com || !offsets) return fail("mmc_upload_chains: null argument");
    CK(cudaSetDevice(h->device));
    const DevConst& dc = h->dc; const size_t C = dc.n_chains; cudaStream_t st = h->stream;
    for (size_t c = 0; c < C; c++)
        for (int r = 0; r < dc.nres; r++)
            if (dc.is_active[r] && (count[c * dc.nres + r] < 0 || count[c * dc.nres + r] > dc.cap[r]))
                return fail("chain %d residue %d: count %d outside [0,%d]", (int)c, r, count[c * dc.nres + r], dc.cap[r]);
    CK(cudaMemcpy2DAsync(h->cs.count, MMC_MAX_RES * sizeof(int), count, dc.nres * sizeof(int), dc.nres * sizeof(int), C, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(h->cs.gcom, com, C * dc.g_mols * 3 * sizeof(double), cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(h->cs.goff, offsets, C * dc.g_atoms * 3 * sizeof(double), cudaMemcpyHostToDevice, st));
    k_rebuild_positions<<<dim3((dc.g_atoms + 255) / 256, (unsigned)C), 256, 0, st>>>(dc, h->cs);
    h->launches++;
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(st));
    for (size_t c = 0; c < C; c++)
        for (int r = 0; r < MMC_MAX_RES; r++)
            h->h_count[c * MMC_MAX_RES + r] = (r < dc.nres && dc.is_active[r]) ? count[c * dc.nres + r] : 0;
    h->pending_chain = -1;
    return 0;
}

extern "C" int mmc_download_chains(mmc_t* h, int* count, double* com, double* offsets, double* energies, mmc_counters* counters) {
    if (!h) return fail("null handle");
    CK(cudaSetDevice(h->device));
    const DevConst& dc = h->dc; const size_t C = dc.n_chains; cudaStream_t st = h->stream;
    if (count) CK(cudaMemcpy2DAsync(count, dc.nres * sizeof(int), h->cs.count, MMC_MAX_RES * sizeof(int), dc.nres * sizeof(int), C, cudaMemcpyDeviceToHost, st));
    if (com) CK(cudaMemcpyAsync(com, h->cs.gcom, C * dc.g_mols * 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (offsets) CK(cudaMemcpyAsync(offsets, h->cs.goff, C * dc.g_atoms * 3 * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (energies) CK(cudaMemcpyAsync(energies, h->cs.energy, C * MMC_NE * sizeof(double), cudaMemcpyDeviceToHost, st));
    if (counters) CK(cudaMemcpyAsync(counters, h->cs.counters, C * sizeof(mmc_counters), cudaMemcpyDeviceToHost, st));
    int* h_flag = (int*)(h->h_out + 15);   // pinned; the last of the 16 result doubles is not used by any energy read-back
    CK(cudaMemcpyAsync(h_flag, h->cs.error_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (count)  // inactive residues: the host already knows them
        for (size_t c = 0; c < C; c++)
            for (int r = 0; r < dc.nres; r++) if (!dc.is_active[r]) count[c * dc.nres + r] = dc.cap[r];
    return report_error_flag(h, *h_flag, true);   // the arrays are filled either way
}

// Per-block snapshot without stalling the chains (write_utils.f90:13-188 runs between blocks in the reference):
// device-to-device copy of the small per-chain state on the handle's stream (the next batch may be queued right behind
// it), then ONE device-to-host copy of the blob into pinned memory on a second stream.
extern "C" int mmc_snapshot_begin(mmc_t* h, int slot) {
    if (!h) return fail("null handle");
    if (slot < 0 || slot > 1) return fail("mmc_snapshot_begin: slot must be 0 or 1");
    CK(cudaSetDevice(h->device));
    const DevConst& dc = h->dc; const size_t C = dc.n_chains;
    if (!h->snap_bytes) {
        const size_t sz[6] = {C * MMC_MAX_RES * sizeof(int), C * dc.g_mols * 3 * sizeof(double), C * dc.g_atoms * 3 * sizeof(double),
                              C * MMC_NE * sizeof(double), C * sizeof(mmc_counters), sizeof(int)};
        size_t o = 0;
        for (int i = 0; i < 6; i++) { h->snap_off[i] = o; o += (sz[i] + 255) & ~(size_t)255; }
        h->snap_bytes = o;
        CK(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    }
    mmc_handle::Snap& s = h->snap[slot];
    if (!s.dev) {
        CK(cudaMalloc((void**)&s.dev, h->snap_bytes));
        CK(cudaMallocHost((void**)&s.host, h->snap_bytes));
        CK(cudaEventCreateWithFlags(&s.taken, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&s.landed, cudaEventDisableTiming));
    }
    if (s.pending) CK(cudaEventSynchronize(s.landed));   // the slot's previous copy-out still reads s.dev
    const void* src[6] = {h->cs.count, h->cs.gcom, h->cs.goff, h->cs.energy, h->cs.counters, h->cs.error_flag};
    const size_t sz[6] = {C * MMC_MAX_RES * sizeof(int), C * dc.g_mols * 3 * sizeof(double), C * dc.g_atoms * 3 * sizeof(double),
                          C * MMC_NE * sizeof(double), C * sizeof(mmc_counters), sizeof(int)};
    for (int i = 0; i < 6; i++)
        if (sz[i]) CK(cudaMemcpyAsync(s.dev + h->snap_off[i], src[i], sz[i], cudaMemcpyDeviceToDevice, h->stream));
    CK(cudaEventRecord(s.taken, h->stream));
    CK(cudaStreamWaitEvent(h->copy_stream, s.taken, 0));
    CK(cudaMemcpyAsync(s.host, s.dev, h->snap_bytes, cudaMemcpyDeviceToHost, h->copy_stream));
    CK(cudaEventRecord(s.landed, h->copy_stream));
    s.pending = true;
    return 0;
}

extern "C" int mmc_snapshot_wait(mmc_t* h, int slot, const int** count, const double** com, const double** offsets,
                                 const double** energies, const mmc_counters** counters) {
    if (!h) return fail("null handle");
    if (slot < 0 || slot > 1 || !h->snap[slot].pending) return fail("mmc_snapshot_wait: no snapshot pending on this slot");
    CK(cudaSetDevice(h->device));
    mmc_handle::Snap& s = h->snap[slot];
    CK(cudaEventSynchronize(s.landed));                  // waits for the copy-out only, not for the handle's stream
    const DevConst& dc = h->dc; const size_t C = dc.n_chains;
    std::vector<int>& cnt = h->snap_count[slot];
    cnt.resize(C * dc.nres);
    const int* raw = (const int*)(s.host + h->snap_off[0]);
    for (size_t c = 0; c < C; c++)
        for (int r = 0; r < dc.nres; r++) cnt[c * dc.nres + r] = dc.is_active[r] ? raw[c * MMC_MAX_RES + r] : dc.cap[r];
    if (count) *count = cnt.data();
    if (com) *com = (const double*)(s.host + h->snap_off[1]);
    if (offsets) *offsets = (const double*)(s.host + h->snap_off[2]);
    if (energies) *energies = (const double*)(s.host + h->snap_off[3]);
    if (counters) *counters = (const mmc_counters*)(s.host + h->snap_off[4]);
    // the error word as it stood when the snapshot was taken (sticky on the device: a later mmc_synchronize reports it again)
    return report_error_flag(h, *(const int*)(s.host + h->snap_off[5]), false);
}

extern "C" int mmc_get_reservoir(mmc_t* h, int c, int* count, double* offsets) {
    CHECK_CHAIN(h, c);
    const DevConst& dc = h->dc;
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    std::vector<int> rc(MMC_MAX_RES, 0);
    CK(cudaMemcpy(rc.data(), h->cs.rcount + (size_t)c * MMC_MAX_RES, MMC_MAX_RES * sizeof(int), cudaMemcpyDeviceToHost));
    size_t dst = 0;
    for (int r = 0; r < dc.nres; r++) {
        if (count) count[r] = rc[r];
        const size_t n = (size_t)rc[r] * dc.natoms[r] * 3;
        if (offsets && n) CK(cudaMemcpy(offsets + dst, h->cs.roff + ((size_t)c * dc.r_atoms + dc.r_atom_base[r]) * 3, n * sizeof(double), cudaMemcpyDeviceToHost));
        dst += n;
    }
    return 0;
}
extern "C" int mmc_get_sk(mmc_t* h, int c, double* sk) {
    CHECK_CHAIN(h, c);
    if (!sk) return fail("mmc_get_sk: null output");
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    int cur = 0;
    CK(cudaMemcpy(&cur, h->cs.cur + c, sizeof(int), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(sk, h->cs.A + ((size_t)c * 2 + cur) * h->dc.nk, (size_t)h->dc.nk * sizeof(double2), cudaMemcpyDeviceToHost));
    return 0;
}

// ---------------------------------------------------------------------------------------------
__global__ void k_store_energy(ChainState cs, int c, const double* res /* lj, coul, recip, -, self, intra */) {
    double* E = cs.energy + c * MMC_NE;
    E[MMC_E_NONCOUL] = res[0]; E[MMC_E_COUL] = res[1]; E[MMC_E_RECIP] = res[2]; E[MMC_E_SELF] = res[4]; E[MMC_E_INTRA] = res[5];
    // energy_utils.f90:32-33
    E[MMC_E_TOTAL] = E[MMC_E_RECIP] + E[MMC_E_NONCOUL] + E[MMC_E_COUL] + E[MMC_E_SELF] + E[MMC_E_INTRA];
}

extern "C" int mmc_system_energy(mmc_t* h, int c, double out[MMC_NE]) {
    CHECK_CHAIN(h, c);
    CK(cudaSetDevice(h->device));
    const DevConst& dc = h->dc; SysScratch& ss = h->ss; cudaStream_t st = h->stream;
    CK(cudaMemsetAsync(ss.tile_min, 0x7f, ss.n_tiles * sizeof(int), st));
    CK(cudaMemsetAsync(ss.tile_max, 0xff, ss.n_tiles * sizeof(int), st));
    k_sys_gather<<<(ss.n + 255) / 256, 256, 0, st>>>(dc, h->cs, ss, c, h->n_fw_mols);
    k_sys_pair<<<dim3(ss.n_tiles, ss.n_tiles), SYS_THREADS, 0, st>>>(dc, ss);
    k_sum_rows<<<1, SYS_THREADS, 0, st>>>(ss.part, (long long)ss.n_tiles * ss.n_tiles, 2, ss.result, 1.0, dc.coul_const);
    k_sys_sk<<<dim3(dc.n_k_tiles, ss.n_super), RECIP_THREADS, h->sk_smem, st>>>(dc, ss, h->sk_chunk);
    k_sys_sk_reduce<<<dc.n_k_tiles, RECIP_THREADS, 0, st>>>(dc, h->cs, ss, c, h->init_sk == MMC_INIT_SK_FULL);
    k_sum_rows<<<1, SYS_THREADS, 0, st>>>(ss.part, dc.n_k_tiles, 2, ss.result + 2, dc.recip_const, 0.0);
    k_sys_self_intra<<<1, SYS_THREADS, 0, st>>>(dc, h->cs, c, ss.result + 4);
    k_store_energy<<<1, 1, 0, st>>>(h->cs, c, ss.result);
    h->launches += 8;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(h->h_out, h->cs.energy + (size_t)c * MMC_NE, MMC_NE * sizeof(double), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (out) memcpy(out, h->h_out, MMC_NE * sizeof(double));
    return 0;
}

// The framework's share of ComputeSystemEnergy -- S_fw(k), framework-framework pair energy, framework self term -- is the same
// for every chain of the handle: computed once, on first use.
static int ensure_framework_terms(mmc_handle* h) {
    if (h->fw_ready) return 0;
    const DevConst& dc = h->dc; cudaStream_t st = h->stream;
    CK(dalloc(h, &h->d_sfw, (size_t)dc.nk)); CK(dalloc(h, &h->d_fwc, 4));   // zero filled: right for a system without framework
    if (dc.n_fw > 0) {
        SysScratch f = h->ss;     // a view of the scratch arrays restricted to the inactive atoms
        f.n = dc.n_fw; f.n_tiles = (f.n + SYS_THREADS - 1) / SYS_THREADS;
        const int n_chunks = (f.n + h->sk_chunk - 1) / h->sk_chunk;
        f.n_super = n_chunks < 128 ? n_chunks : 128;
        CK(cudaMemsetAsync(f.tile_min, 0x7f, f.n_tiles * sizeof(int), st));
        CK(cudaMemsetAsync(f.tile_max, 0xff, f.n_tiles * sizeof(int), st));
        k_sys_gather<<<(f.n + 255) / 256, 256, 0, st>>>(dc, h->cs, f, 0, h->n_fw_mols);
        k_sys_pair<<<dim3(f.n_tiles, f.n_tiles), SYS_THREADS, 0, st>>>(dc, f);
        k_sum_rows<<<1, SYS_THREADS, 0, st>>>(f.part, (long long)f.n_tiles * f.n_tiles, 2, h->d_fwc, 1.0, dc.coul_const);
        k_sys_sk<<<dim3(dc.n_k_tiles, f.n_super), RECIP_THREADS, h->sk_smem, st>>>(dc, f, h->sk_chunk);
        k_sys_sk_store<<<dc.n_k_tiles, RECIP_THREADS, 0, st>>>(dc, f, h->d_sfw);
        k_sys_fw_self<<<1, SYS_THREADS, 0, st>>>(dc, h->d_fwc + 2);
        h->launches += 6;
        CK(cudaGetLastError());
    }
    h->fw_ready = true;
    return 0;
}

// ComputeSystemEnergy for EVERY chain of the handle in one launch set and one synchronisation: the framework terms come
// from ensure_framework_terms, each chain adds its guests (A_chain = S_fw + S_guests).  out: [n_chains][MMC_NE] or NULL.
extern "C" int mmc_system_energy_all(mmc_t* h, double* out) {
    if (!h) return fail("null handle");
    CK(cudaSetDevice(h->device));
    const DevConst& dc = h->dc; cudaStream_t st = h->stream; const int C = dc.n_chains;
    if (ensure_framework_terms(h)) return 1;
    SysAll sa;
    sa.n_other_tiles = (dc.n_fw + dc.g_atoms + SYS_THREADS - 1) / SYS_THREADS;
    if (!h->d_pair_part) {
        CK(dalloc(h, &h->d_pair_part, (size_t)C * sa.n_other_tiles * 2));
        CK(dalloc(h, &h->d_recip_part, (size_t)C * dc.n_k_tiles));
        CK(dalloc(h, &h->d_eall, (size_t)C * MMC_NE));
    }
    sa.sfw = h->d_sfw; sa.fwc = h->d_fwc; sa.pair_part = h->d_pair_part; sa.recip_part = h->d_recip_part; sa.out = h->d_eall;
    sa.n_fw_mols = h->n_fw_mols; sa.store_sk = h->init_sk == MMC_INIT_SK_FULL; sa.chain0 = 0;
    const size_t smem = (size_t)h->sk_chunk * 3 * dc.kstride * sizeof(double2);
    for (int c0 = 0; c0 < C; c0 += 32768) {   // gridDim.y limit
        const int nc = std::min(32768, C - c0);
        SysAll sb = sa; sb.chain0 = c0;
        sb.pair_part = sa.pair_part + (size_t)c0 * sa.n_other_tiles * 2; sb.recip_part = sa.recip_part + (size_t)c0 * dc.n_k_tiles;
        sb.out = sa.out + (size_t)c0 * MMC_NE;
        k_sys_pair_guests<<<dim3(sa.n_other_tiles, nc), SYS_THREADS, 0, st>>>(dc, h->cs, sb);
        k_sys_sk_guests<<<dim3(dc.n_k_tiles, nc), RECIP_THREADS, smem, st>>>(dc, h->cs, sb, h->sk_chunk);
        k_sys_finish_all<<<nc, SYS_THREADS, 0, st>>>(dc, h->cs, sb, dc.n_k_tiles);
        h->launches += 3;
    }
    CK(cudaGetLastError());
    if (out) CK(cudaMemcpyAsync(out, h->d_eall, (size_t)C * MMC_NE * sizeof(double), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return 0;
}

// ---------------------------------------------------------------------------------------------
static void launch_sweeps(mmc_handle* h, cudaStream_t st, int chain) {
    k_sweep<<<dim3(h->dc.sweep_parts, 1), SWEEP_THREADS, h->sweep_smem, st>>>(h->dc, h->cs, chain);
    h->launches += 1;
}

// ---- strict mode ------------------------------------------------------------------------------
static int check_trial(mmc_handle* h, int c, int kind, int res, int mol, const double* com_new, const double* offs_new, HostTrial& ht,
                       const char* who, int res2 = -1) {
    CHECK_CHAIN(h, c);
    const DevConst& dc = h->dc;
    if (h->pending_chain >= 0) return fail("%s: a trial is already pending on chain %d (call mmc_trial_commit first)", who, h->pending_chain);
    if (h->pending_chain == -2) return fail("%s: a sharded step is open (call mmc_shard_step_end first)", who);
    if (kind < MMC_CREATION || kind > MMC_SWAP) return fail("Unknown move_type %d", kind);  // monte_carlo_utils.f90:218
    if (res < 0 || res >= dc.nres || !dc.is_active[res]) return fail("%s: residue %d is not an active residue type", who, res);
    const int cnt = h->h_count[(size_t)c * MMC_MAX_RES + res];
    int nat_new = dc.natoms[res];
    if (kind == MMC_SWAP) {
        if (res2 < 0 || res2 >= dc.nres || !dc.is_active[res2] || res2 == res) return fail("%s: swap needs a second, different active residue type (got %d)", who, res2);
        const int cnt2 = h->h_count[(size_t)c * MMC_MAX_RES + res2];
        if (cnt2 >= dc.cap[res2]) return fail("molecule index %d exceeds the maximum allowed = %d", cnt2 + 1, dc.cap[res2]);
        nat_new = dc.natoms[res2];
    }
    if (kind == MMC_CREATION) {
        if (mol != cnt) return fail("%s: creation must target slot %d (got %d)", who, cnt, mol);
        if (mol >= dc.cap[res]) return fail("molecule index %d exceeds the maximum allowed = %d", mol + 1, dc.cap[res]);  // CheckMoleculeIndex
    } else if (mol < 0 || mol >= cnt) return fail("%s: molecule %d out of range [0,%d)", who, mol, cnt);
    if (kind != MMC_DELETION && (!com_new || !offs_new)) return fail("%s: proposed geometry missing", who);
    memset(&ht, 0, sizeof ht);
    ht.chain = c; ht.kind = kind; ht.res = res; ht.mol = mol; ht.res2 = kind == MMC_SWAP ? res2 : res;
    if (kind != MMC_DELETION) {
        for (int d = 0; d < 3; d++) ht.com[d] = com_new[d];
        for (int a = 0; a < nat_new; a++) for (int d = 0; d < 3; d++) ht.off[a][d] = offs_new[a * 3 + d];
    }
    h->pending_res2 = ht.res2;
    return 0;
}

// One launch per strict-mode trial (k_strict_trial): the result comes back through host-mapped pinned memory and the host
// spins on its sequence number -- no D2H copy, no stream synchronisation.  Round 1 needed k_set_trial + k_sweep + k_finalize +
// a copy + a synchronisation (37 us per c4_large_k trial).
static int strict_trial(mmc_handle* h, const HostTrial& ht, int decide, double u, const char* who) {
    const DevConst& dc = h->dc;
    if (dc.shard_count > 1) return fail("%s: this handle is shard %d of %d; use mmc_shard_trial_begin/_end", who, dc.shard_rank, dc.shard_count);
    CK(cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    const unsigned long long seq = ++h->strict_seq;
    k_strict_trial<<<dc.sweep_parts, SWEEP_THREADS, h->sweep_smem, st>>>(dc, h->cs, ht, decide, u, h->d_ticket, h->d_sout, seq);
    h->launches += 1;
    CK(cudaGetLastError());
    volatile StrictOut* o = h->h_sout;
    for (unsigned long long spins = 0; o->seq != seq; spins++) {
        if ((spins & 0xfffffull) == 0xfffffull) {      // every ~1e6 spins: has the stream died?
            const cudaError_t e = cudaStreamQuery(st);
            if (e != cudaSuccess && e != cudaErrorNotReady) return fail("%s: kernel failed: %s", who, cudaGetErrorString(e));
            if (e == cudaSuccess && o->seq != seq) return fail("%s: the trial kernel finished without reporting a result", who);
        }
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    return 0;
}

static int trial_energy_impl(mmc_handle* h, const HostTrial& ht, double old_e[MMC_NE], double new_e[MMC_NE], const char* who) {
    if (strict_trial(h, ht, 0, 0.0, who)) return 1;
    if (old_e) memcpy(old_e, (const void*)h->h_sout->old_e, MMC_NE * sizeof(double));
    if (new_e) memcpy(new_e, (const void*)h->h_sout->new_e, MMC_NE * sizeof(double));
    h->pending_chain = ht.chain; h->pending_kind = ht.kind; h->pending_res = ht.res;
    return 0;
}

// SaveMoleculeState + ComputeOldEnergy + ComputeNewEnergy + mc_acceptance_probability + the accept test + Accept* / Reject* of
// ONE move in one launch: `u` is the uniform the move routine draws for its accept test (translation.f90:66, rotation.f90:62,
// create_molecule.f90:77, delete_molecule.f90:80), handed over with the proposal.  accepted: 1 / 0.
extern "C" int mmc_trial_step(mmc_t* h, int c, int kind, int res, int mol, int res_new, const double* com_new, const double* offs_new, double u,
                              double old_e[MMC_NE], double new_e[MMC_NE], double* prob, int* accepted) {
    HostTrial ht;
    if (check_trial(h, c, kind, res, mol, com_new, offs_new, ht, "mmc_trial_step", kind == MMC_SWAP ? res_new : -1)) return 1;
    if (strict_trial(h, ht, 1, u, "mmc_trial_step")) return 1;
    const StrictOut* o = h->h_sout;
    if (old_e) memcpy(old_e, o->old_e, MMC_NE * sizeof(double));
    if (new_e) memcpy(new_e, o->new_e, MMC_NE * sizeof(double));
    if (prob) *prob = o->prob;
    if (accepted) *accepted = o->accepted;
    if (o->accepted == 1) {
        int* cnt = h->h_count.data() + (size_t)c * MMC_MAX_RES;
        if (kind == MMC_CREATION) cnt[res]++;
        if (kind == MMC_DELETION) cnt[res]--;
        if (kind == MMC_SWAP) { cnt[res]--; cnt[ht.res2]++; }
    }
    h->pending_chain = -1;
    return 0;
}

extern "C" int mmc_trial_energy(mmc_t* h, int c, int kind, int res, int mol, const double* com_new, const double* offs_new,
                                double old_e[MMC_NE], double new_e[MMC_NE]) {
    HostTrial ht;
    if (kind == MMC_SWAP) return fail("mmc_trial_energy: a swap names two residue types; use mmc_trial_energy_swap");
    if (check_trial(h, c, kind, res, mol, com_new, offs_new, ht, "mmc_trial_energy")) return 1;
    return trial_energy_impl(h, ht, old_e, new_e, "mmc_trial_energy");
}
extern "C" int mmc_trial_energy_swap(mmc_t* h, int c, int res_old, int mol, int res_new, const double* com_new, const double* offs_new,
                                     double old_e[MMC_NE], double new_e[MMC_NE]) {
    HostTrial ht;
    if (check_trial(h, c, MMC_SWAP, res_old, mol, com_new, offs_new, ht, "mmc_trial_energy_swap", res_new)) return 1;
    return trial_energy_impl(h, ht, old_e, new_e, "mmc_trial_energy_swap");
}

// k-sharded strict mode: the same trial on every rank; begin leaves this rank's six partial sums in dev_sums (device,
// enqueued on mmc_stream), the caller all-reduces them in place, end assembles old/new.  mmc_trial_commit as usual.
extern "C" int mmc_shard_trial_begin(mmc_t* h, int c, int kind, int res, int mol, const double* com_new, const double* offs_new,
                                     double* dev_sums) {
    HostTrial ht;
    if (!dev_sums) return fail("mmc_shard_trial_begin: dev_sums must be a device buffer of 6 doubles");
    if (kind == MMC_SWAP) return fail("mmc_shard_trial_begin: swap trials are not available through the sharded strict entry point");
    if (check_trial(h, c, kind, res, mol, com_new, offs_new, ht, "mmc_shard_trial_begin")) return 1;
    CK(cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    k_set_trial<<<1, TAIL_THREADS, 0, st>>>(h->dc, h->cs, ht);
    launch_sweeps(h, st, c);
    k_part_sum<<<1, TAIL_THREADS, 0, st>>>(h->dc, h->cs, c, dev_sums);
    h->launches += 2;
    CK(cudaGetLastError());
    h->pending_chain = c; h->pending_kind = kind; h->pending_res = res;
    return 0;
}
extern "C" int mmc_shard_trial_end(mmc_t* h, int c, const double* dev_sums, double old_e[MMC_NE], double new_e[MMC_NE]) {
    CHECK_CHAIN(h, c);
    if (!dev_sums) return fail("mmc_shard_trial_end: dev_sums missing");
    if (h->pending_chain != c) return fail("mmc_shard_trial_end: no trial pending on chain %d", c);
    CK(cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    k_finish<<<1, 32, 0, st>>>(h->dc, h->cs, h->d_ctl, c, dev_sums, 0, h->d_out);
    h->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(h->h_out, h->d_out, (2 * MMC_NE + 1) * sizeof(double), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (old_e) memcpy(old_e, h->h_out, MMC_NE * sizeof(double));
    if (new_e) memcpy(new_e, h->h_out + MMC_NE, MMC_NE * sizeof(double));
    return 0;
}

extern "C" int mmc_trial_commit(mmc_t* h, int c, int accept) {
    CHECK_CHAIN(h, c);
    if (h->pending_chain != c) return fail("mmc_trial_commit: no trial pending on chain %d", c);
    CK(cudaSetDevice(h->device));
    k_commit<<<1, 32, 0, h->stream>>>(h->dc, h->cs, c, accept ? 1 : 0);
    h->launches++;
    CK(cudaGetLastError());
    if (accept) {
        if (h->pending_kind == MMC_CREATION) h->h_count[(size_t)c * MMC_MAX_RES + h->pending_res]++;
        if (h->pending_kind == MMC_DELETION) h->h_count[(size_t)c * MMC_MAX_RES + h->pending_res]--;
        if (h->pending_kind == MMC_SWAP) { h->h_count[(size_t)c * MMC_MAX_RES + h->pending_res]--; h->h_count[(size_t)c * MMC_MAX_RES + h->pending_res2]++; }
    }
    h->pending_chain = -1;
    return 0;
}

// ---- batched mode -------------------------------------------------------------------------------
#ifndef GRAPH_CHUNK
#define GRAPH_CHUNK 128
#endif

// one launch of the chain kernel: every chain advances n steps.  philox: the random numbers are device Philox draws (not a
// recorded tape): with few chains the cluster-per-chain kernel is used then
static int launch_chain_run(mmc_handle* h, cudaStream_t st, int n, int phases, bool philox = true) {
    const int C = h->dc.n_chains;
    if (h->cluster_p > 1 && philox) {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(C * h->cluster_p)); cfg.blockDim = dim3(CLUSTER_THREADS);
        cfg.dynamicSmemBytes = h->cluster_smem; cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = (unsigned)h->cluster_p; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        const BatchCtl* ctl = h->d_ctl;
        cudaError_t e = h->dc.tri ? cudaLaunchKernelEx(&cfg, k_chain_run_cl<CLUSTER_THREADS, true>, h->dc, h->cs, ctl, 0, n, phases)
                                  : cudaLaunchKernelEx(&cfg, k_chain_run_cl<CLUSTER_THREADS, false>, h->dc, h->cs, ctl, 0, n, phases);
        h->launches++;
        if (e != cudaSuccess) return fail("cluster launch of k_chain_run_cl (P = %d) failed: %s", h->cluster_p, cudaGetErrorString(e));
        return 0;
    }
    if (h->dc.tri) {   // the tilted-cell build of the same kernel
        if (h->chain_threads == 256) k_chain_run<256, 4, true><<<C, 256, h->chain_smem, st>>>(h->dc, h->cs, h->d_ctl, 0, n, phases);
        else if (h->chain_threads == 512) k_chain_run<512, 2, true><<<C, 512, h->chain_smem, st>>>(h->dc, h->cs, h->d_ctl, 0, n, phases);
        else k_chain_run<1024, 1, true><<<C, 1024, h->chain_smem, st>>>(h->dc, h->cs, h->d_ctl, 0, n, phases);
    }
    else if (h->chain_threads == 256) k_chain_run<256, 4, false><<<C, 256, h->chain_smem, st>>>(h->dc, h->cs, h->d_ctl, 0, n, phases);
    else if (h->chain_threads == 512) k_chain_run<512, 2, false><<<C, 512, h->chain_smem, st>>>(h->dc, h->cs, h->d_ctl, 0, n, phases);
    else k_chain_run<1024, 1, false><<<C, 1024, h->chain_smem, st>>>(h->dc, h->cs, h->d_ctl, 0, n, phases);
    h->launches++;
    return 0;
}
static int enqueue_steps(mmc_handle* h, cudaStream_t st, int n, bool philox = true) {
    if (launch_chain_run(h, st, n, 3, philox)) return 1;
    k_advance<<<1, 1, 0, st>>>(h->d_ctl, n);
    h->launches++;
    return 0;
}

static int enqueue_batch(mmc_handle* h, int64_t n_steps, uint64_t seed, const double* d_tape, int64_t tape_stride,
                         mmc_move_record* d_recs) {
    cudaStream_t st = h->stream;
    if (h->dc.shard_count > 1) return fail("mmc_run_batch: this handle is shard %d of %d; use mmc_shard_step_begin/_end", h->dc.shard_rank, h->dc.shard_count);
    if (h->pending_chain == -2) return fail("mmc_run_batch: a sharded step is open");
    BatchCtl ctl;
    ctl.rng.tape = d_tape; ctl.rng.tape_stride = tape_stride; ctl.rng.seed = seed; ctl.rng.step = h->steps_done;
    ctl.recs = d_recs; ctl.rec_stride = n_steps; ctl.rec_base = 0;
    CK(cudaMemcpyAsync(h->d_ctl, &ctl, sizeof ctl, cudaMemcpyHostToDevice, st));  // pageable source: copied before return
    CK(cudaMemsetAsync(h->cs.tape_pos, 0, h->dc.n_chains * sizeof(long long), st));
    // cs.error_flag is NOT cleared here: it stays set until a host entry point reports it (report_error_flag)
    int64_t done = 0;
    // the cached graph holds the kernel chosen for Philox runs; a tape replay on a cluster handle runs the one-CTA kernel directly
    const bool philox = d_tape == nullptr, use_graph = philox || h->cluster_p <= 1;
    if (!use_graph)
        for (; n_steps - done >= GRAPH_CHUNK; done += GRAPH_CHUNK) if (enqueue_steps(h, st, GRAPH_CHUNK, false)) return 1;
    if (use_graph && n_steps >= GRAPH_CHUNK) {
        if (!h->graph) {
            cudaGraph_t g;
            const int64_t l0 = h->launches;
            CK(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
            const int cap_rc = enqueue_steps(h, st, GRAPH_CHUNK);
            CK(cudaStreamEndCapture(st, &g));
            if (cap_rc) return 1;
            h->launches = l0;
            CK(cudaGraphInstantiate(&h->graph, g, 0));
            CK(cudaGraphDestroy(g));
            h->graph_steps = GRAPH_CHUNK;
        }
        while (n_steps - done >= GRAPH_CHUNK) {
            CK(cudaGraphLaunch(h->graph, st));
            h->launches += 2;
            done += GRAPH_CHUNK;
        }
    }
    if (n_steps - done > 0 && enqueue_steps(h, st, (int)(n_steps - done), philox)) return 1;
    CK(cudaGetLastError());
    h->steps_done += (unsigned long long)n_steps;
    h->pending_chain = -1;
    return 0;
}

static int refresh_counts(mmc_handle* h) {
    CK(cudaMemcpy(h->h_count.data(), h->cs.count, h->h_count.size() * sizeof(int), cudaMemcpyDeviceToHost));
    int flag = 0;
    CK(cudaMemcpy(&flag, h->cs.error_flag, sizeof(int), cudaMemcpyDeviceToHost));
    return report_error_flag(h, flag, true);
}

extern "C" int mmc_run_batch_async(mmc_t* h, int64_t n_steps, uint64_t seed) {
    if (!h) return fail("null handle");
    if (n_steps < 1) return fail("mmc_run_batch: n_steps must be >= 1");
    CK(cudaSetDevice(h->device));
    return enqueue_batch(h, n_steps, seed, nullptr, 0, nullptr);
}

extern "C" int mmc_profile_kernels(mmc_t* h, int n_steps, uint64_t seed, double ms[3], int64_t n[3]) {
    if (!h || !ms || !n) return fail("mmc_profile_kernels: null argument");
    if (n_steps < 1 || n_steps > 4096) return fail("mmc_profile_kernels: n_steps must be in [1,4096]");
    CK(cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    BatchCtl ctl;
    ctl.rng.tape = nullptr; ctl.rng.tape_stride = 0; ctl.rng.seed = seed; ctl.rng.step = h->steps_done;
    ctl.recs = nullptr; ctl.rec_stride = n_steps; ctl.rec_base = 0;
    CK(cudaMemcpyAsync(h->d_ctl, &ctl, sizeof ctl, cudaMemcpyHostToDevice, st));
    cudaEvent_t ev[4];
    for (auto& e : ev) CK(cudaEventCreate(&e));
    // the same n_steps proposals three times: real-space phase only, reciprocal phase only (neither commits,
    // so the chains do not move), then the full step, which does commit
    const int phases[3] = {1, 2, 3};
    mmc_counters* d_save = nullptr;   // the single-phase passes must not count as trial moves
    CK(cudaMalloc((void**)&d_save, h->dc.n_chains * sizeof(mmc_counters)));
    CK(cudaMemcpyAsync(d_save, h->cs.counters, h->dc.n_chains * sizeof(mmc_counters), cudaMemcpyDeviceToDevice, st));
    for (int k = 0; k < 3; k++) {
        if (k == 2) CK(cudaMemcpyAsync(h->cs.counters, d_save, h->dc.n_chains * sizeof(mmc_counters), cudaMemcpyDeviceToDevice, st));
        CK(cudaEventRecord(ev[k], st));
        if (launch_chain_run(h, st, n_steps, phases[k])) return 1;
    }
    CK(cudaEventRecord(ev[3], st));
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(st));
    for (int k = 0; k < 3; k++) { float t = 0; CK(cudaEventElapsedTime(&t, ev[k], ev[k + 1])); ms[k] = t; n[k] = 1; }
    for (auto& e : ev) cudaEventDestroy(e);
    cudaFree(d_save);
    h->steps_done += (unsigned long long)n_steps;
    h->pending_chain = -1;
    return refresh_counts(h);
}

extern "C" int mmc_synchronize(mmc_t* h) {
    if (!h) return fail("null handle");
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    return refresh_counts(h);
}

extern "C" int mmc_run_batch(mmc_t* h, int64_t n_steps, uint64_t seed, const double* tape, int64_t tape_stride,
                             mmc_move_record* records, mmc_counters* counters) {
    if (!h) return fail("null handle");
    if (n_steps < 1) return fail("mmc_run_batch: n_steps must be >= 1");
    CK(cudaSetDevice(h->device));
    const size_t C = h->dc.n_chains;
    cudaStream_t st = h->stream;
    const double* d_tape = nullptr;
    if (tape) {
        if (tape_stride < 1) return fail("mmc_run_batch: tape_stride must be >= 1");
        const size_t n = C * (size_t)tape_stride;
        if (n > h->tape_cap) { if (h->d_tape) cudaFree(h->d_tape); h->d_tape = nullptr; h->tape_cap = 0; CK(cudaMalloc((void**)&h->d_tape, n * sizeof(double))); h->tape_cap = n; }
        CK(cudaMemcpyAsync(h->d_tape, tape, n * sizeof(double), cudaMemcpyHostToDevice, st));
        d_tape = h->d_tape;
    }
    mmc_move_record* d_recs = nullptr;
    if (records) {
        const size_t n = C * (size_t)n_steps;
        if (n > h->recs_cap) { if (h->d_recs) cudaFree(h->d_recs); h->d_recs = nullptr; h->recs_cap = 0; CK(cudaMalloc((void**)&h->d_recs, n * sizeof(mmc_move_record))); h->recs_cap = n; }
        d_recs = h->d_recs;
    }
    if (enqueue_batch(h, n_steps, seed, d_tape, tape_stride, d_recs)) return 1;
    if (records) CK(cudaMemcpyAsync(records, d_recs, C * (size_t)n_steps * sizeof(mmc_move_record), cudaMemcpyDeviceToHost, st));
    if (counters) CK(cudaMemcpyAsync(counters, h->cs.counters, C * sizeof(mmc_counters), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(h->h_tape_pos, h->cs.tape_pos, C * sizeof(long long), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return refresh_counts(h);
}

// ---- k-sharded mode (one handle per GPU, dc.shard_rank / dc.shard_count) ---------------------------------------
static int shard_check(mmc_handle* h, const void* dev_sums, const char* who) {
    if (!h) return fail("null handle");
    if (!dev_sums) return fail("%s: dev_sums must be a device buffer of 6 doubles per trial", who);
    return 0;
}
static int shard_ctl(mmc_handle* h, uint64_t seed, int with_records) {
    const size_t C = h->dc.n_chains;
    if (with_records && C > h->recs_cap) {
        if (h->d_recs) cudaFree(h->d_recs);
        h->d_recs = nullptr; h->recs_cap = 0;
        CK(cudaMalloc((void**)&h->d_recs, C * sizeof(mmc_move_record))); h->recs_cap = C;
    }
    BatchCtl ctl;
    ctl.rng.tape = nullptr; ctl.rng.tape_stride = 0; ctl.rng.seed = seed; ctl.rng.step = h->steps_done;
    ctl.recs = with_records ? h->d_recs : nullptr; ctl.rec_stride = 1; ctl.rec_base = 0;
    CK(cudaMemcpyAsync(h->d_ctl, &ctl, sizeof ctl, cudaMemcpyHostToDevice, h->stream));
    return 0;
}

extern "C" int mmc_shard_range(const mmc_t* h, int* k_first, int* k_count) {
    if (!h) return fail("null handle");
    const DevConst& dc = h->dc;
    const int ny = 2 * dc.kmax[1] + 1, nrows = (dc.kmax[0] + 1) * ny;
    const int n_parts = dc.shard_count * dc.sweep_parts, rpp = (nrows + n_parts - 1) / n_parts;
    const int r0 = std::min(nrows, dc.shard_rank * dc.sweep_parts * rpp), r1 = std::min(nrows, (dc.shard_rank + 1) * dc.sweep_parts * rpp);
    if (k_first) *k_first = h->h_row_kstart[r0];
    if (k_count) *k_count = h->h_row_kstart[r1] - h->h_row_kstart[r0];
    return 0;
}

extern "C" int mmc_shard_step_begin(mmc_t* h, uint64_t seed, double* dev_sums) {
    if (shard_check(h, dev_sums, "mmc_shard_step_begin")) return 1;
    if (h->pending_chain >= 0) return fail("mmc_shard_step_begin: a strict-mode trial is pending on chain %d", h->pending_chain);
    CK(cudaSetDevice(h->device));
    const DevConst& dc = h->dc; const int C = dc.n_chains; cudaStream_t st = h->stream;
    if (shard_ctl(h, seed, 1)) return 1;
    k_propose_all<<<C, 32, 0, st>>>(dc, h->cs, h->d_ctl, 0);
    k_sweep<<<dim3(dc.sweep_parts, C), SWEEP_THREADS, h->sweep_smem, st>>>(dc, h->cs, 0);
    k_part_sum<<<C, TAIL_THREADS, 0, st>>>(dc, h->cs, 0, dev_sums);
    h->launches += 3;
    CK(cudaGetLastError());
    h->pending_chain = -2;   // a sharded step is open
    return 0;
}

extern "C" int mmc_shard_step_end(mmc_t* h, const double* dev_sums, mmc_move_record* records) {
    if (shard_check(h, dev_sums, "mmc_shard_step_end")) return 1;
    if (h->pending_chain != -2) return fail("mmc_shard_step_end: no sharded step is open");
    CK(cudaSetDevice(h->device));
    const DevConst& dc = h->dc; const int C = dc.n_chains; cudaStream_t st = h->stream;
    k_finish<<<C, 32, 0, st>>>(dc, h->cs, h->d_ctl, 0, dev_sums, 1, nullptr);
    h->launches++;
    CK(cudaGetLastError());
    h->steps_done += 1;
    h->pending_chain = -1;
    if (records) {
        CK(cudaMemcpyAsync(records, h->d_recs, (size_t)C * sizeof(mmc_move_record), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        return refresh_counts(h);
    }
    return 0;
}

extern "C" int64_t mmc_tape_used(mmc_t* h, int c) {
    if (!h || c < 0 || c >= h->dc.n_chains) return -1;
    return (int64_t)h->h_tape_pos[c];
}

// ---- debug / test entry points -----------------------------------------------------------------
extern "C" int mmc_dbg_pair_singlemol(mmc_t* h, int c, int res, int mol, double* e_nc, double* e_c) {
    // a deletion trial requests exactly one sweep at the current geometry with the molecule excluded
    double o[MMC_NE], n[MMC_NE];
    if (mmc_trial_energy(h, c, MMC_DELETION, res, mol, nullptr, nullptr, o, n)) return 1;
    if (mmc_trial_commit(h, c, 0)) return 1;
    if (e_nc) *e_nc = o[MMC_E_NONCOUL];
    if (e_c) *e_c = o[MMC_E_COUL];
    return 0;
}
extern "C" int mmc_dbg_self_singlemol(mmc_t* h, int res, double* e_self) {
    if (!h) return fail("null handle");
    if (res < 0 || res >= h->dc.nres || !h->dc.is_active[res]) return fail("residue %d is not active", res);
    if (e_self) *e_self = h->dc.self_e[res];
    return 0;
}
extern "C" int mmc_dbg_intra_singlemol(mmc_t* h, int c, int res, int mol, double* e_intra) {
    double o[MMC_NE], n[MMC_NE];
    if (mmc_trial_energy(h, c, MMC_DELETION, res, mol, nullptr, nullptr, o, n)) return 1;
    if (mmc_trial_commit(h, c, 0)) return 1;
    if (e_intra) *e_intra = o[MMC_E_INTRA];
    return 0;
}
// Measurement helper for bench.py's strict_mode object: n_moves translation moves of chain c driven from the host EXACTLY as
// the Fortran Translation routine drives them (translation.f90:36-112: pick a molecule, displace its centre of mass by
// (u - 1/2) step, ApplyPBC, energies, Metropolis test, accept / reject), through the public entry points -- fused = 1:
// mmc_trial_step, one launch per move; fused = 0: mmc_trial_energy + host decision + mmc_trial_commit.  The host keeps its
// own mirror of the coordinates like the Fortran driver does.  Returns wall seconds and the number of accepted moves.
extern "C" int mmc_dbg_strict_loop(mmc_t* h, int c, int n_moves, uint64_t seed, int fused, double* seconds, int64_t* n_accepted) {
    CHECK_CHAIN(h, c);
    if (n_moves < 1 || !seconds) return fail("mmc_dbg_strict_loop: bad arguments");
    const DevConst& dc = h->dc;
    if (dc.tri) return fail("mmc_dbg_strict_loop: orthorhombic cells only");
    if (dc.n_active < 1) return fail("mmc_dbg_strict_loop: no active residue");
    const int r = dc.active_list[0], na = dc.natoms[r];
    std::vector<int> cnt(dc.nres);
    size_t nm = 0, nat = 0, mo = 0, ao = 0;
    for (int q = 0; q < dc.nres; q++) { const size_t cap = dc.is_active[q] ? dc.cap[q] : dc.cap[q]; if (q < r) { } nm += cap; nat += cap * dc.natoms[q]; }
    std::vector<double> com(3 * nm + 3), off(3 * nat + 3);
    if (download_chain(h, c, cnt.data(), com.data(), off.data())) return 1;
    for (int q = 0; q < r; q++) { mo += cnt[q]; ao += (size_t)cnt[q] * dc.natoms[q]; }
    if (cnt[r] < 1) return fail("mmc_dbg_strict_loop: chain %d has no molecule of the first active type", c);
    double tstep = 1.0, temp = 300.0;
    CK(cudaMemcpy(&tstep, h->cs.tstep + c, sizeof(double), cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(&temp, h->cs.temp + c, sizeof(double), cudaMemcpyDeviceToHost));
    uint64_t x = seed * 6364136223846793005ull + 1442695040888963407ull;
    auto uni = [&]() { x = x * 6364136223846793005ull + 1442695040888963407ull; return (double)(x >> 11) * (1.0 / 9007199254740992.0); };
    int64_t nacc = 0;
    double o[MMC_NE], nw[MMC_NE];
    CK(cudaStreamSynchronize(h->stream));
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < n_moves; i++) {
        const int m = std::min((int)(uni() * cnt[r]), cnt[r] - 1);
        double* cm = com.data() + (mo + m) * 3; const double* of = off.data() + (ao + (size_t)m * na) * 3;
        double cn[3];
        for (int d = 0; d < 3; d++) {
            double v = cm[d] + (uni() - 0.5) * tstep - dc.lo[d];
            v -= std::floor(v / dc.L[d]) * dc.L[d];
            cn[d] = dc.lo[d] + v;
        }
        int acc = 0;
        if (fused) {
            double p;
            if (mmc_trial_step(h, c, MMC_TRANSLATION, r, m, -1, cn, of, uni(), o, nw, &p, &acc)) return 1;
        } else {
            if (mmc_trial_energy(h, c, MMC_TRANSLATION, r, m, cn, of, o, nw)) return 1;
            const double p = std::fmin(1.0, std::exp(-(nw[MMC_E_TOTAL] - o[MMC_E_TOTAL]) / temp));
            acc = uni() <= p;
            if (mmc_trial_commit(h, c, acc)) return 1;
        }
        if (acc) { nacc++; for (int d = 0; d < 3; d++) cm[d] = cn[d]; }
    }
    CK(cudaStreamSynchronize(h->stream));
    *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (n_accepted) *n_accepted = nacc;
    return 0;
}

// test support: overwrite a chain's tracked energies and / or its tracked A(k) (resynchronising a long parity run with
// its CPU checker after an accept/reject flip; the reference has no counterpart)
extern "C" int mmc_dbg_set_tracked(mmc_t* h, int c, const double* energies, const double* sk) {
    CHECK_CHAIN(h, c);
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    if (energies) CK(cudaMemcpy(h->cs.energy + (size_t)c * MMC_NE, energies, MMC_NE * sizeof(double), cudaMemcpyHostToDevice));
    if (sk) {
        int cur = 0;
        CK(cudaMemcpy(&cur, h->cs.cur + c, sizeof(int), cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(h->cs.A + ((size_t)c * 2 + cur) * h->dc.nk, sk, (size_t)h->dc.nk * sizeof(double2), cudaMemcpyHostToDevice));
    }
    return 0;
}
static int run_elementwise(mmc_handle* h, int which, const double* in, int n, double* out) {
    if (!h || !in || !out || n < 1) return fail("bad arguments");
    CK(cudaSetDevice(h->device));
    double *di = nullptr, *dout = nullptr;
    CK(cudaMalloc((void**)&di, n * sizeof(double))); CK(cudaMalloc((void**)&dout, n * sizeof(double)));
    CK(cudaMemcpyAsync(di, in, n * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (which < 0) k_erfc_test<<<(n + 255) / 256, 256, 0, h->stream>>>(h->dc, di, n, dout);
    else k_table_lookup<<<(n + 255) / 256, 256, 0, h->stream>>>(h->dc, which, di, n, dout);
    h->launches++;
    cudaError_t e = cudaStreamSynchronize(h->stream);
    if (e == cudaSuccess) e = cudaMemcpy(out, dout, n * sizeof(double), cudaMemcpyDeviceToHost);
    cudaFree(di); cudaFree(dout);
    if (e != cudaSuccess) return fail("elementwise kernel failed: %s", cudaGetErrorString(e));
    return 0;
}
extern "C" int mmc_dbg_table_lookup(mmc_t* h, int which, const double* r, int n, double* out) {
    if (which < 0 || which > 2) return fail("which must be 0, 1 or 2");
    return run_elementwise(h, which, r, n, out);
}
extern "C" int mmc_dbg_erfc(mmc_t* h, const double* x, int n, double* out) { return run_elementwise(h, -1, x, n, out); }

// ---- allocation-free variant of mmc_dbg_erfc (no counterpart in the reference) ----
// For callers that may not allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc): the same kernel as mmc_dbg_erfc (k_erfc_test: the per-move sweep's erfc, Coulomb-kernel table + slow path), but
// the device staging scratch that run_elementwise cudaMallocs and frees on every call lives in ONE file-scope device array of
// fixed size.  Every GPU that loads the library holds these 2 x MMC_NOALLOC_MAX_N doubles (1 MiB) until the process exits.
//   * g_noalloc_scratch below is, in HOST code, only a host-side stand-in for the symbol -- NOT a device address.  The device
//     address is taken once per device, with that device current, by cudaGetSymbolAddress and cached; only that pointer is
//     ever handed to a kernel or a copy.
//   * All calls on one GPU share the array and it keeps the previous call's contents.  Nothing here has to start at zero:
//     the first n input slots are overwritten by the upload and the first n output slots by the kernel before they are read,
//     and nothing beyond n is read.  Calls that OVERLAP on the GPU (two handles / two host threads on one device) race on it:
//     the caller serialises them.  This function itself returns only after its own work on the stream has drained.
static __device__ __align__(256) double g_noalloc_scratch[2 * MMC_NOALLOC_MAX_N];   // [0, MAX): input staging, [MAX, 2 MAX): output
#define MMC_NOALLOC_MAX_DEVICES 64
static double* g_noalloc_ptr[MMC_NOALLOC_MAX_DEVICES] = {nullptr};   // device address of g_noalloc_scratch on device d
static std::mutex g_noalloc_mutex;

// set-up, once per device; the caller has made `device` current
static int noalloc_scratch_for(int device, double** out) {
    if (device < 0 || device >= MMC_NOALLOC_MAX_DEVICES) return fail("mmc_erfc_noalloc: device ordinal %d not supported (max %d)", device, MMC_NOALLOC_MAX_DEVICES - 1);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device]) {
        double* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
        if (e != cudaSuccess || !ptr) return fail("mmc_erfc_noalloc: cudaGetSymbolAddress failed on device %d: %s", device, cudaGetErrorString(e));
        g_noalloc_ptr[device] = ptr;
    }
    *out = g_noalloc_ptr[device];
    return 0;
}

extern "C" int mmc_erfc_noalloc(mmc_t* h, const double* x, int n, double* out) {
    if (!h || !x || !out) return fail("mmc_erfc_noalloc: null argument");
    if (n < 1) return fail("mmc_erfc_noalloc: n must be >= 1");
    if (n > MMC_NOALLOC_MAX_N) return fail("mmc_erfc_noalloc: n = %d exceeds the fixed scratch of MMC_NOALLOC_MAX_N = %d values; split the call or use mmc_dbg_erfc", n, MMC_NOALLOC_MAX_N);
    CK(cudaSetDevice(h->device));
    double* scratch = nullptr;
    if (noalloc_scratch_for(h->device, &scratch)) return 1;
    double* d_in = scratch; double* d_out = scratch + MMC_NOALLOC_MAX_N;
    cudaStream_t st = h->stream;
    CK(cudaMemcpyAsync(d_in, x, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, st));
    k_erfc_test<<<(n + 255) / 256, 256, 0, st>>>(h->dc, d_in, n, d_out);
    h->launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out, d_out, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    return 0;
}
extern "C" void* mmc_stream(mmc_t* h) { return h ? (void*)h->stream : nullptr; }
extern "C" int64_t mmc_launch_count(const mmc_t* h) { return h ? h->launches : -1; }
extern "C" int mmc_chain_launch_config(const mmc_t* h, int* threads, int* cluster) {
    if (!h) return fail("null handle");
    if (threads) *threads = h->cluster_p > 1 ? CLUSTER_THREADS : h->chain_threads;
    if (cluster) *cluster = h->cluster_p > 1 ? h->cluster_p : 1;
    return 0;
}
extern "C" int mmc_pair_work(mmc_t* h, int64_t out[2]) {
    if (!h || !out) return fail("mmc_pair_work: null argument");
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    unsigned long long v[2];
    CK(cudaMemcpy(v, h->cs.pair_count, sizeof v, cudaMemcpyDeviceToHost));
    out[0] = (int64_t)v[0]; out[1] = (int64_t)v[1];
    return 0;
}

// -DMMC_PHASE_TIMING builds: summed clock64 cycles of thread 0 of every CTA per phase since mmc_init (slots: propose, stage,
// cull + real-space sweep, block sum, reciprocal sweep, block sum, finalize + commit, trials); zeros in production builds
extern "C" int mmc_dbg_phase_cycles(mmc_t* h, int64_t out[8]) {
    if (!h || !out) return fail("mmc_dbg_phase_cycles: null argument");
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    unsigned long long v[8];
    CK(cudaMemcpy(v, h->cs.pair_count + 2, sizeof v, cudaMemcpyDeviceToHost));
    for (int i = 0; i < 8; i++) out[i] = (int64_t)v[i];
    return 0;
}

// the framework-energy cache of one chain (ChainState.fw_pair): out[2 * slot + {0, 1}] = {LJ, Coulomb in reduced units} of the
// molecule in device slot `slot` (slots of mmc_chain_capacity's g_mols), NaN where nothing is cached
extern "C" int mmc_dbg_fw_cache(mmc_t* h, int chain, double* out) {
    CHECK_CHAIN(h, chain);
    if (!out) return fail("mmc_dbg_fw_cache: null argument");
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaMemcpy(out, h->cs.fw_pair + (size_t)chain * h->dc.g_mols, (size_t)h->dc.g_mols * sizeof(double2), cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int mmc_measure_fp64_peak(int device, double* tflops) {
    if (!tflops) return fail("null output");
    CK(cudaSetDevice(device));
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, device));
    const int blocks = p.multiProcessorCount * 8, threads = 256, iters = 20000;
    double* d = nullptr; CK(cudaMalloc((void**)&d, (size_t)blocks * threads * sizeof(double)));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    k_fp64_peak<<<blocks, threads>>>(d, 1000);
    double best = 0.0;
    for (int rep = 0; rep < 5; rep++) {
        CK(cudaEventRecord(e0));
        k_fp64_peak<<<blocks, threads>>>(d, iters);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
        const double tf = 2.0 * 8.0 * iters * (double)blocks * threads / (ms * 1e-3) / 1e12;
        if (tf > best) best = tf;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
    *tflops = best;
    return 0;
}
