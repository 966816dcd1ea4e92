/*
 * maniac_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C, glibc libm, sequential sums in the reference's loop order) of
 * MANIAC's per-move energy path, used as the parity checker for the CUDA library and as the
 * timed CPU baseline ("port") in bench.py.  Nothing in the product path may link or call it.
 *
 * Reference: maniac-mc/maniac-mc @ 961f5c9 (pure Fortran; cannot be compiled in this image:
 * no gfortran/flang).  Every function cites the reference file:line it follows
 * (paths relative to the reference's src/).
 *
 * Pinning: single-point routines are pinned by the reference's two energy KATs
 * (tests/integration/energy/{LJ-gas,H2O-gas}/run-test.sh, +-0.01 kcal/mol), see
 * tests/test_oracle_golden.py.  The per-move routines and accept/reject sequences have NO
 * golden vector in the reference ("parity unpinned" for those); they are checked
 * through consistency identities instead (tracked totals == fresh system energy, etc.).
 *
 * Indices are 0-based here (Fortran is 1-based).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "../include/maniac_gpu.h" /* mmc_desc, mmc_move_record, mmc_counters: shared data shapes only */

/* constants.f90:7-20 */
static const double PI = 3.14159265358979323846;
#define TWOPI (2.0 * PI)
static const double EPS0_INV_eVA = 14.40198;
static const double KB_eVK = 8.6173852e-5;
static const double ERR = 1.0e-10;
/* parameters.f90:15-22,41 */
static const double TARGET_ACCEPTANCE = 0.40, TOL_ACCEPTANCE = 0.05;
static const double MIN_TRANSLATION_STEP = 1.0e-3, MAX_TRANSLATION_STEP = 3.0;
static const double MIN_ROTATION_ANGLE = 1.0e-3, MAX_ROTATION_ANGLE = 0.78;
static const double PROB_CREATE_DELETE = 0.5;
#define MIN_TRIALS_FOR_RECALIBRATION 500
#define TABULATED_POINTS 5000

typedef struct { double re, im; } cplx;
static inline cplx cmul(cplx a, cplx b) { cplx r = {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; return r; }

typedef struct { double *x, *f; double dx; int n; int initialized; } orc_table; /* simulation_state.f90:194-204 */

/* random source: recorded tape, or Philox4x32-10 with the same (seed, chain, step, draw)
 * indexing as the device batched mode */
typedef struct {
    const double* tape; int64_t pos, len;
    uint64_t seed; uint32_t chain; uint64_t step; uint32_t draw;
} orc_rng;

typedef struct orc {
    int nres, max_atoms, ntypes;
    int *natoms, *is_active, *count, *cap;
    double *charges; int *types;
    double *sigma, *epsilon;
    double **com;  /* com[r][3*m + d]                      primary%mol_com(d,r,m)        */
    double **off;  /* off[r][(m*natoms[r] + a)*3 + d]      primary%site_offset(d,r,m,a)  */
    double lo[3], L[3], volume;
    double tilt[3]; int tri; double half_min2;   /* xy, xz, yz; any non-zero; (min(L)/2)^2 */
    double alpha, rc, T; double* fugacity;
    double translation_step, rotation_step, p_translation, p_rotation, p_swap;
    int kmax[3], nk; int *kx, *ky, *kz; double *ksq_mag, *form, *W;
    cplx *A, *A_old;                     /* ewald%recip_amplitude(_old) */
    cplx **px, **py, **pz;               /* ewald%phase_factor_{x,y,z}(r, m, a, -kmax:kmax) */
    cplx *px_old, *py_old, *pz_old;      /* ewald%phase_factor_*_old(a, -kmax:kmax) */
    double energy[MMC_NE];               /* type(energy_type) energy */
    double com_old[3]; double* off_old;  /* res%mol_com_old, res%site_offset_old */
    int init_sk, deletion_mode, use_table, recalibrate;
    int has_reservoir; int *rsv_count, *rsv_cap; double** rsv_off;   /* reservoir%num_residues, reservoir%site_offset */
    orc_table erfc_r, r6, r12;
    mmc_counters cnt;
    orc_rng rng;
} orc_t;

#define PH(o, P, ax, r, m, a, k) ((o)->P[r][((size_t)((m) * (o)->natoms[r] + (a))) * (2 * (o)->kmax[ax] + 1) + ((k) + (o)->kmax[ax])])
#define PHOLD(o, P, ax, a, k) ((o)->P[(size_t)(a) * (2 * (o)->kmax[ax] + 1) + ((k) + (o)->kmax[ax])])

/* gfortran modulo(a,p) for reals: a - floor(a/p)*p evaluated as fmod + sign fix-up */
static inline double f_modulo(double a, double p) {
    double r = fmod(a, p);
    if (r != 0.0 && ((r < 0.0) != (p < 0.0))) r += p;
    return r;
}

/* ---------------- Philox4x32-10 (Salmon et al. 2011), shared definition with the device ------ */
static void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int i = 0; i < 10; i++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1, n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
void orc_philox(uint32_t ctr[4], uint32_t k0, uint32_t k1) { philox4x32_10(ctr, k0, k1); }

/* uniform in [0,1): draw d of step s of chain c = 53 top bits of words (2*(d&1), 2*(d&1)+1)
 * of Philox(counter = (s_lo, s_hi, c, d>>1), key = seed) */
double orc_philox_uniform(uint64_t seed, uint32_t chain, uint64_t step, uint32_t draw) {
    uint32_t c[4] = {(uint32_t)step, (uint32_t)(step >> 32), chain, draw >> 1};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    uint32_t lo = c[2 * (draw & 1)], hi = c[2 * (draw & 1) + 1];
    uint64_t bits = (((uint64_t)hi << 32) | lo) >> 11;
    return (double)bits * (1.0 / 9007199254740992.0);
}

static double rand_uniform(orc_t* o) { /* random_utils.f90:13 */
    orc_rng* g = &o->rng;
    if (g->tape) {
        if (g->pos >= g->len) { fprintf(stderr, "oracle: RNG tape exhausted\n"); abort(); }
        return g->tape[g->pos++];
    }
    return orc_philox_uniform(g->seed, g->chain, g->step, g->draw++);
}

/* ---------------- Ewald set-up --------------------------------------------------------------- */

/* ComputeEwaldParameters + ComputeFourierIndices, prepare_utils.f90:169-214
 * (CheckValidReciprocalVector ewald_kvectors.f90:199, NormalizedKSquared :105) */
static int valid_k(int kx, int ky, int kz, const int kmax[3], double* ksq) {
    double a = (double)kx / (double)kmax[0], b = (double)ky / (double)kmax[1], c = (double)kz / (double)kmax[2];
    double k2 = a * a + b * b + c * c;
    if (ksq) *ksq = k2;
    return (fabs(k2) >= ERR) && (k2 <= 1.0);
}
void orc_ewald_setup(double rc, double tol, const double L[3], double* alpha, int kmax[3], int* nk) {
    tol = fmin(fabs(tol), 0.5);                                   /* ClampTolerance :158 */
    double s = sqrt(fabs(log(tol * rc)));                         /* screening_factor :171 */
    double a = sqrt(fabs(log(tol * rc * s))) / rc;                /* alpha :172 */
    double p = sqrt(-log(tol * rc * (2.0 * s * a) * (2.0 * s * a))); /* fourier_precision :174 */
    for (int d = 0; d < 3; d++) kmax[d] = (int)lround(0.25 + L[d] * a * p / PI); /* nint :195 */
    int count = 0;
    for (int kx = 0; kx <= kmax[0]; kx++)
        for (int ky = -kmax[1]; ky <= kmax[1]; ky++)
            for (int kz = -kmax[2]; kz <= kmax[2]; kz++) {
                if (kx == 0 && ky == 0 && kz == 0) continue;
                if (valid_k(kx, ky, kz, kmax, NULL)) count++;
            }
    *alpha = a; *nk = count;
}

/* PrecomputeValidReciprocalVectors ewald_kvectors.f90:44, ComputeCartesianKSquared :134,
 * ComputeSymmetryFormFactor :167.  Orthorhombic: reciprocal = diag(1/L). */
static void precompute_kvectors(orc_t* o) {
    int n = 0;
    for (int pass = 0; pass < 2; pass++) {
        n = 0;
        for (int kx = 0; kx <= o->kmax[0]; kx++)
            for (int ky = -o->kmax[1]; ky <= o->kmax[1]; ky++)
                for (int kz = -o->kmax[2]; kz <= o->kmax[2]; kz++) {
                    if (kx == 0 && ky == 0 && kz == 0) continue;
                    if (!valid_k(kx, ky, kz, o->kmax, NULL)) continue;
                    if (pass == 1) {
                        double v[3] = {(double)kx * (TWOPI * (1.0 / o->L[0])), (double)ky * (TWOPI * (1.0 / o->L[1])),
                                       (double)kz * (TWOPI * (1.0 / o->L[2]))};
                        if (o->tri) { /* k = 2 pi (a b c)^-T n for a = (lx,0,0), b = (xy,ly,0), c = (xz,yz,lz) */
                            const double gx = (double)kx * (1.0 / o->L[0]);
                            const double gy = ((double)ky - o->tilt[0] * gx) * (1.0 / o->L[1]);
                            const double gz = ((double)kz - o->tilt[1] * gx - o->tilt[2] * gy) * (1.0 / o->L[2]);
                            v[0] = TWOPI * gx; v[1] = TWOPI * gy; v[2] = TWOPI * gz;
                        }
                        o->kx[n] = kx; o->ky[n] = ky; o->kz[n] = kz;
                        o->ksq_mag[n] = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
                        o->form[n] = (kx == 0) ? 1.0 : 2.0;
                    }
                    n++;
                }
        if (pass == 0) {
            o->nk = n;
            o->kx = malloc(sizeof(int) * n); o->ky = malloc(sizeof(int) * n); o->kz = malloc(sizeof(int) * n);
            o->ksq_mag = malloc(sizeof(double) * n); o->form = malloc(sizeof(double) * n); o->W = calloc(n, sizeof(double));
        }
    }
}

/* ComputeReciprocalWeights ewald_kvectors.f90:225 */
static void reciprocal_weights(orc_t* o) {
    double a2 = o->alpha * o->alpha;
    for (int i = 0; i < o->nk; i++) o->W[i] = exp(-o->ksq_mag[i] / (4.0 * a2)) / o->ksq_mag[i];
}

/* ---------------- tables, tabulated_utils.f90 -------------------------------------------------- */
static void table_alloc(orc_table* t, double rcut) {
    t->x = malloc(sizeof(double) * (TABULATED_POINTS + 1)); t->f = malloc(sizeof(double) * (TABULATED_POINTS + 1));
    t->n = TABULATED_POINTS; t->dx = rcut / (double)TABULATED_POINTS;
}
static void init_table_erfc_r(orc_table* t, double alpha, double rcut) { /* :19 */
    table_alloc(t, rcut);
    for (int i = 0; i <= TABULATED_POINTS; i++) {
        double r = i * t->dx; t->x[i] = r;
        t->f[i] = (r < ERR) ? 2.0 * alpha / sqrt(PI) : erfc(alpha * r) / r;
    }
    t->initialized = 1;
}
static void init_table_rpower(orc_table* t, double rcut, int power) { /* :52 */
    table_alloc(t, rcut);
    for (int i = 0; i <= TABULATED_POINTS; i++) {
        double r = i * t->dx; t->x[i] = r;
        if (r < ERR) t->f[i] = 0.0;
        else { double p = 1.0; for (int k = 0; k < power; k++) p *= r; t->f[i] = p; }
    }
    t->initialized = 1;
}
static double lookup_tabulated(const orc_table* t, double r) { /* :87 */
    if (r <= 0.0) return t->f[0];
    if (r >= t->x[t->n]) return 0.0;
    int i = (int)(r / t->dx);
    double f1 = t->f[i], f2 = t->f[i + 1];
    double w = (r - t->x[i]) / t->dx;
    return (1.0 - w) * f1 + w * f2;
}
double orc_table_lookup(orc_t* o, int which, double r) {
    return lookup_tabulated(which == 0 ? &o->erfc_r : which == 1 ? &o->r6 : &o->r12, r);
}

/* ---------------- construction ------------------------------------------------------------------ */
orc_t* orc_create(const mmc_desc* d) {
    orc_t* o = calloc(1, sizeof(orc_t));
    o->nres = d->nres; o->max_atoms = d->max_atoms; o->ntypes = d->ntypes;
    size_t nr = d->nres;
    o->natoms = malloc(nr * sizeof(int)); o->is_active = malloc(nr * sizeof(int));
    o->count = malloc(nr * sizeof(int)); o->cap = malloc(nr * sizeof(int));
    o->fugacity = malloc(nr * sizeof(double));
    memcpy(o->natoms, d->natoms, nr * sizeof(int)); memcpy(o->is_active, d->is_active, nr * sizeof(int));
    memcpy(o->count, d->count, nr * sizeof(int));
    for (size_t r = 0; r < nr; r++) {
        o->cap[r] = d->capacity ? d->capacity[r] : d->count[r];
        if (o->cap[r] < d->count[r]) o->cap[r] = d->count[r];
        if (o->cap[r] < 1) o->cap[r] = 1;
        o->fugacity[r] = d->fugacity ? d->fugacity[r] : 0.0;
    }
    o->charges = malloc(nr * d->max_atoms * sizeof(double)); memcpy(o->charges, d->charges, nr * d->max_atoms * sizeof(double));
    o->types = malloc(nr * d->max_atoms * sizeof(int)); memcpy(o->types, d->types, nr * d->max_atoms * sizeof(int));
    size_t nt2 = (size_t)d->ntypes * d->ntypes;
    o->sigma = malloc(nt2 * sizeof(double)); memcpy(o->sigma, d->sigma, nt2 * sizeof(double));
    o->epsilon = malloc(nt2 * sizeof(double)); memcpy(o->epsilon, d->epsilon, nt2 * sizeof(double));
    for (int k = 0; k < 3; k++) { o->lo[k] = d->box_lo[k]; o->L[k] = d->box_len[k]; o->kmax[k] = d->kmax[k]; }
    o->volume = o->L[0] * o->L[1] * o->L[2];
    for (int k = 0; k < 3; k++) o->tilt[k] = d->tilt[k];
    o->tri = o->tilt[0] != 0.0 || o->tilt[1] != 0.0 || o->tilt[2] != 0.0;
    { double m = fmin(o->L[0], fmin(o->L[1], o->L[2])); o->half_min2 = 0.25 * m * m; }
    o->alpha = d->alpha; o->rc = d->cutoff; o->T = d->temperature;
    o->translation_step = d->translation_step; o->rotation_step = d->rotation_step;
    o->p_translation = d->p_translation; o->p_rotation = d->p_rotation; o->p_swap = d->p_swap;
    o->init_sk = d->init_sk; o->deletion_mode = d->deletion_mode; o->use_table = d->use_table;
    /* coordinates */
    o->com = malloc(nr * sizeof(double*)); o->off = malloc(nr * sizeof(double*));
    o->px = malloc(nr * sizeof(cplx*)); o->py = malloc(nr * sizeof(cplx*)); o->pz = malloc(nr * sizeof(cplx*));
    size_t mo = 0, ao = 0;
    for (size_t r = 0; r < nr; r++) {
        size_t cap = o->cap[r], na = o->natoms[r];
        o->com[r] = calloc(cap * 3, sizeof(double)); o->off[r] = calloc(cap * na * 3, sizeof(double));
        memcpy(o->com[r], d->com + 3 * mo, sizeof(double) * 3 * o->count[r]);
        memcpy(o->off[r], d->offsets + 3 * ao, sizeof(double) * 3 * o->count[r] * na);
        mo += o->count[r]; ao += (size_t)o->count[r] * na;
        o->px[r] = calloc(cap * na * (2 * o->kmax[0] + 1), sizeof(cplx));
        o->py[r] = calloc(cap * na * (2 * o->kmax[1] + 1), sizeof(cplx));
        o->pz[r] = calloc(cap * na * (2 * o->kmax[2] + 1), sizeof(cplx));
    }
    o->px_old = calloc((size_t)o->max_atoms * (2 * o->kmax[0] + 1), sizeof(cplx));
    o->py_old = calloc((size_t)o->max_atoms * (2 * o->kmax[1] + 1), sizeof(cplx));
    o->pz_old = calloc((size_t)o->max_atoms * (2 * o->kmax[2] + 1), sizeof(cplx));
    o->off_old = calloc((size_t)o->max_atoms * 3, sizeof(double));
    /* reservoir box: only its site offsets matter to the path (data_parser.f90:60-62, create_molecule.f90:185-193) */
    o->has_reservoir = d->reservoir_count != NULL && d->reservoir_offsets != NULL;
    o->rsv_count = calloc(nr, sizeof(int)); o->rsv_cap = calloc(nr, sizeof(int)); o->rsv_off = calloc(nr, sizeof(double*));
    if (o->has_reservoir) {
        size_t ro = 0;
        for (size_t r = 0; r < nr; r++) {
            size_t na = o->natoms[r];
            o->rsv_count[r] = d->reservoir_count[r]; o->rsv_cap[r] = d->reservoir_count[r] + o->cap[r] + 1;
            o->rsv_off[r] = calloc((size_t)o->rsv_cap[r] * na * 3, sizeof(double));
            memcpy(o->rsv_off[r], d->reservoir_offsets + 3 * ro, sizeof(double) * 3 * na * o->rsv_count[r]);
            ro += (size_t)o->rsv_count[r] * na;
        }
    }
    precompute_kvectors(o);
    reciprocal_weights(o);
    o->A = calloc(o->nk, sizeof(cplx)); o->A_old = calloc(o->nk, sizeof(cplx)); /* Q1: de-facto zero */
    init_table_erfc_r(&o->erfc_r, o->alpha, o->rc); /* PrecomputeTable tabulated_utils.f90:10 */
    init_table_rpower(&o->r6, o->rc, 6);
    init_table_rpower(&o->r12, o->rc, 12);
    return o;
}

void orc_destroy(orc_t* o) {
    if (!o) return;
    for (int r = 0; r < o->nres; r++) { free(o->com[r]); free(o->off[r]); free(o->px[r]); free(o->py[r]); free(o->pz[r]); }
    free(o->com); free(o->off); free(o->px); free(o->py); free(o->pz);
    free(o->px_old); free(o->py_old); free(o->pz_old); free(o->off_old);
    free(o->natoms); free(o->is_active); free(o->count); free(o->cap); free(o->fugacity);
    free(o->charges); free(o->types); free(o->sigma); free(o->epsilon);
    free(o->kx); free(o->ky); free(o->kz); free(o->ksq_mag); free(o->form); free(o->W);
    free(o->A); free(o->A_old);
    for (int r = 0; r < o->nres; r++) free(o->rsv_off[r]);
    free(o->rsv_count); free(o->rsv_cap); free(o->rsv_off);
    free(o->erfc_r.x); free(o->erfc_r.f); free(o->r6.x); free(o->r6.f); free(o->r12.x); free(o->r12.f);
    free(o);
}

int orc_num_kvectors(orc_t* o) { return o->nk; }
void orc_get_kvectors(orc_t* o, int* kx, int* ky, int* kz, double* W, double* form) {
    for (int i = 0; i < o->nk; i++) {
        if (kx) kx[i] = o->kx[i];
        if (ky) ky[i] = o->ky[i];
        if (kz) kz[i] = o->kz[i];
        if (W) W[i] = o->W[i];
        if (form) form[i] = o->form[i];
    }
}
void orc_set_params(orc_t* o, const double* fugacity, double T) {
    if (fugacity) memcpy(o->fugacity, fugacity, sizeof(double) * o->nres);
    o->T = T;
}
void orc_set_recalibrate(orc_t* o, int flag) { o->recalibrate = flag; }
void orc_get_step_sizes(orc_t* o, double* t, double* r) { *t = o->translation_step; *r = o->rotation_step; }
void orc_get_state(orc_t* o, int* count, double* com, double* offsets) {
    size_t mo = 0, ao = 0;
    for (int r = 0; r < o->nres; r++) {
        size_t na = o->natoms[r];
        if (count) count[r] = o->count[r];
        if (com) memcpy(com + 3 * mo, o->com[r], sizeof(double) * 3 * o->count[r]);
        if (offsets) memcpy(offsets + 3 * ao, o->off[r], sizeof(double) * 3 * o->count[r] * na);
        mo += o->count[r]; ao += (size_t)o->count[r] * na;
    }
}
void orc_get_reservoir(orc_t* o, int* count, double* offsets) {
    size_t ao = 0;
    for (int r = 0; r < o->nres; r++) {
        size_t na = o->natoms[r];
        if (count) count[r] = o->rsv_count[r];
        if (offsets && o->rsv_off[r]) memcpy(offsets + 3 * ao, o->rsv_off[r], sizeof(double) * 3 * o->rsv_count[r] * na);
        ao += (size_t)o->rsv_count[r] * na;
    }
}
void orc_get_sk(orc_t* o, double* sk) { memcpy(sk, o->A, sizeof(cplx) * o->nk); }
void orc_get_energy(orc_t* o, double* e) { memcpy(e, o->energy, sizeof(double) * MMC_NE); }
void orc_get_counters(orc_t* o, mmc_counters* c) { *c = o->cnt; }

/* ---------------- geometry ---------------------------------------------------------------------- */

/* Triclinic nearest image of a separation (ComputeDistance's triclinic branch, geometry_utils.f90:397-411, with the
 * lattice vectors a = (lx,0,0), b = (xy,ly,0), c = (xz,yz,lz) as the shifts - the reference's own shifts are the columns
 * of a matrix whose ROWS are a, b, c: Q8).  The cascade z -> y -> x lands in the brick [-l/2, l/2]^3, a fundamental
 * domain of the lattice; a brick point shorter than half the shortest edge is the nearest image, any other one gets
 * the reference's 27-image search around it.  Same arithmetic as min_image_tri in csrc/common.cuh. */
static void min_image_tri(const orc_t* o, double* d) {
    const double big = 6755399441055744.0; /* 1.5 * 2^52: fma(x, 1/l, big) - big = x/l rounded to an integer */
    double n = fma(d[2], 1.0 / o->L[2], big) - big;
    d[2] = fma(-n, o->L[2], d[2]); d[1] = fma(-n, o->tilt[2], d[1]); d[0] = fma(-n, o->tilt[1], d[0]);
    n = fma(d[1], 1.0 / o->L[1], big) - big;
    d[1] = fma(-n, o->L[1], d[1]); d[0] = fma(-n, o->tilt[0], d[0]);
    n = fma(d[0], 1.0 / o->L[0], big) - big;
    d[0] = fma(-n, o->L[0], d[0]);
    double best = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
    if (best < o->half_min2) return;
    const double x0 = d[0], y0 = d[1], z0 = d[2];
    for (int sa = -1; sa <= 1; sa++)
        for (int sb = -1; sb <= 1; sb++)
            for (int sc = -1; sc <= 1; sc++) {
                const double x = x0 + sa * o->L[0] + sb * o->tilt[0] + sc * o->tilt[1];
                const double y = y0 + sb * o->L[1] + sc * o->tilt[2];
                const double z = z0 + sc * o->L[2];
                const double r2 = x * x + y * y + z * z;
                if (r2 < best) { best = r2; d[0] = x; d[1] = y; d[2] = z; }
            }
}
/* fractional coordinates s = (a b c)^-1 p of a triclinic cell (back substitution) */
static void frac_tri(const orc_t* o, const double* p, double* s) {
    s[2] = (1.0 / o->L[2]) * p[2];
    s[1] = (1.0 / o->L[1]) * (p[1] - o->tilt[2] * s[2]);
    s[0] = (1.0 / o->L[0]) * (p[0] - o->tilt[0] * s[1] - o->tilt[1] * s[2]);
}
static void cart_tri(const orc_t* o, const double* s, double* p) {
    p[0] = o->L[0] * s[0] + o->tilt[0] * s[1] + o->tilt[1] * s[2];
    p[1] = o->L[1] * s[1] + o->tilt[2] * s[2];
    p[2] = o->L[2] * s[2];
}

/* ComputeDistance geometry_utils.f90:359-411 */
static double compute_distance(const orc_t* o, int r1, int m1, int a1, int r2, int m2, int a2) {
    double delta[3], s = 0.0;
    const double* c1 = o->com[r1] + 3 * m1; const double* o1 = o->off[r1] + ((size_t)m1 * o->natoms[r1] + a1) * 3;
    const double* c2 = o->com[r2] + 3 * m2; const double* o2 = o->off[r2] + ((size_t)m2 * o->natoms[r2] + a2) * 3;
    if (o->tri) {
        for (int d = 0; d < 3; d++) delta[d] = (c2[d] + o2[d]) - (c1[d] + o1[d]);
        min_image_tri(o, delta);
        return sqrt(delta[0] * delta[0] + delta[1] * delta[1] + delta[2] * delta[2]);
    }
    for (int d = 0; d < 3; d++) {
        delta[d] = (c2[d] + o2[d]) - (c1[d] + o1[d]);
        delta[d] = f_modulo(delta[d] + 0.5 * o->L[d], o->L[d]) - 0.5 * o->L[d];
        s += delta[d] * delta[d];
    }
    return sqrt(s);
}

/* ApplyPBC geometry_utils.f90:167-219; triclinic: s = (a b c)^-1 (r - r0), s mod 1, r = r0 + (a b c) s */
static void apply_pbc(const orc_t* o, double* pos) {
    if (o->tri) {
        double rel[3] = {pos[0] - o->lo[0], pos[1] - o->lo[1], pos[2] - o->lo[2]}, s[3];
        frac_tri(o, rel, s);
        for (int d = 0; d < 3; d++) { s[d] -= floor(s[d]); if (s[d] >= 1.0) s[d] = 0.0; }
        cart_tri(o, s, rel);
        for (int d = 0; d < 3; d++) pos[d] = o->lo[d] + rel[d];
        return;
    }
    for (int d = 0; d < 3; d++) pos[d] = o->lo[d] + f_modulo(pos[d] - o->lo[d], o->L[d]);
}

static inline double sigma_of(const orc_t* o, int r1, int a1, int r2, int a2) {
    return o->sigma[(size_t)(o->types[r1 * o->max_atoms + a1] - 1) * o->ntypes + (o->types[r2 * o->max_atoms + a2] - 1)];
}
static inline double epsilon_of(const orc_t* o, int r1, int a1, int r2, int a2) {
    return o->epsilon[(size_t)(o->types[r1 * o->max_atoms + a1] - 1) * o->ntypes + (o->types[r2 * o->max_atoms + a2] - 1)];
}
static inline double pow6(double x) { double x2 = x * x; return x2 * x2 * x2; }

/* LennardJonesEnergy energy_utils.f90:192-218 */
static double lennard_jones(const orc_t* o, double r, double sigma, double epsilon) {
    if (r >= o->rc) return 0.0;
    double r6, r12;
    if (o->use_table) {
        r6 = pow6(sigma) / lookup_tabulated(&o->r6, r);
        r12 = pow6(sigma) * pow6(sigma) / lookup_tabulated(&o->r12, r);
    } else { r6 = pow6(sigma / r); r12 = r6 * r6; }
    return 4.0 * epsilon * (r12 - r6);
}
/* CoulombEnergy energy_utils.f90:225-256 */
static double coulomb_energy(const orc_t* o, double r, double q1, double q2) {
    if (fabs(q1) < ERR || fabs(q2) < ERR) return 0.0;
    if (r < ERR) return 0.0;
    if (o->use_table) return q1 * q2 * lookup_tabulated(&o->erfc_r, r);
    return q1 * q2 * erfc(o->alpha * r) / r;
}

/* SingleMolPairwiseEnergy energy_utils.f90:121-187 */
static void single_mol_pairwise(const orc_t* o, int r1, int m1, double* e_nc, double* e_c) {
    double enc = 0.0, ec = 0.0;
    for (int a1 = 0; a1 < o->natoms[r1]; a1++)
        for (int r2 = 0; r2 < o->nres; r2++)
            for (int m2 = 0; m2 < o->count[r2]; m2++) {
                if (m1 == m2 && r1 == r2) continue;
                if (r2 < r1 || (r2 == r1 && m2 <= m1)) continue;
                for (int a2 = 0; a2 < o->natoms[r2]; a2++) {
                    double d = compute_distance(o, r1, m1, a1, r2, m2, a2);
                    enc += lennard_jones(o, d, sigma_of(o, r1, a1, r2, a2), epsilon_of(o, r1, a1, r2, a2));
                    ec += coulomb_energy(o, d, o->charges[r1 * o->max_atoms + a1], o->charges[r2 * o->max_atoms + a2]);
                }
            }
    *e_nc = enc; *e_c = ec * EPS0_INV_eVA / KB_eVK;
}
/* ComputePairwiseEnergy energy_utils.f90:83-115 */
static void pairwise_energy(orc_t* o) {
    o->energy[MMC_E_NONCOUL] = 0.0; o->energy[MMC_E_COUL] = 0.0;
    for (int r = 0; r < o->nres; r++)
        for (int m = 0; m < o->count[r]; m++) {
            double a, b; single_mol_pairwise(o, r, m, &a, &b);
            o->energy[MMC_E_NONCOUL] += a; o->energy[MMC_E_COUL] += b;
        }
}

/* ComputePairInteractionEnergy_singlemol energy_utils.f90:374-442 (no table branch, Q12;
 * no Coulomb cut-off, Q3) */
void orc_pair_singlemol(const orc_t* o, int r1, int m1, double* e_nc, double* e_c) {
    double enc = 0.0, ec = 0.0;
    for (int a1 = 0; a1 < o->natoms[r1]; a1++)
        for (int r2 = 0; r2 < o->nres; r2++)
            for (int m2 = 0; m2 < o->count[r2]; m2++) {
                if (m1 == m2 && r1 == r2) continue;
                for (int a2 = 0; a2 < o->natoms[r2]; a2++) {
                    double d = compute_distance(o, r1, m1, a1, r2, m2, a2);
                    if (d < o->rc) {
                        double r6 = pow6(sigma_of(o, r1, a1, r2, a2) / d);
                        double r12 = r6 * r6;
                        enc += 4.0 * epsilon_of(o, r1, a1, r2, a2) * (r12 - r6);
                    }
                    double q1 = o->charges[r1 * o->max_atoms + a1], q2 = o->charges[r2 * o->max_atoms + a2];
                    if (fabs(q1) < ERR || fabs(q2) < ERR) continue;
                    ec += q1 * q2 * erfc(o->alpha * d) / d;
                }
            }
    *e_nc = enc; *e_c = ec * EPS0_INV_eVA / KB_eVK;
}

/* SingleMolEwaldSelf energy_utils.f90:340 == ComputeEwaldSelfInteractionSingleMol ewald_energy.f90:308 */
double orc_self_singlemol(const orc_t* o, int r) {
    double e = 0.0;
    for (int a = 0; a < o->natoms[r]; a++) {
        double q = o->charges[r * o->max_atoms + a];
        if (fabs(q) < ERR) continue;
        e = e - o->alpha / sqrt(PI) * (q * q);
    }
    return e * EPS0_INV_eVA / KB_eVK;
}
/* ComputeIntraResidueRealCoulombEnergySingleMol ewald_energy.f90:371-411 */
double orc_intra_singlemol(const orc_t* o, int r, int m) {
    double e = 0.0;
    for (int a1 = 0; a1 < o->natoms[r] - 1; a1++) {
        double q1 = o->charges[r * o->max_atoms + a1];
        for (int a2 = a1 + 1; a2 < o->natoms[r]; a2++) {
            double q2 = o->charges[r * o->max_atoms + a2];
            double d = compute_distance(o, r, m, a1, r, m, a2);
            if (d > ERR) e += q1 * q2 * (erfc(o->alpha * d) - 1.0) / d;
        }
    }
    return e * EPS0_INV_eVA / KB_eVK;
}

/* ---------------- phase factors, ewald_phase.f90 ------------------------------------------------- */

/* ComputePhaseFactors1D :90-109: direct cos/sin(k*theta), negative k by conjugation */
static void phase_factors_1d(cplx* axis_mid /* points at k=0 */, double theta, int kmax) {
    for (int k = 0; k <= kmax; k++) {
        cplx e = {cos(k * theta), sin(k * theta)};
        axis_mid[k] = e;
        if (k != 0) { axis_mid[-k].re = e.re; axis_mid[-k].im = -e.im; }
    }
}
/* SingleMolFourierTerms :383-420 with ComputeAtomPhase :41-63 (orthorhombic reciprocal = diag(1/L)) */
void orc_single_mol_fourier(orc_t* o, int r, int m) {
    for (int a = 0; a < o->natoms[r]; a++) {
        double pos[3], th[3];
        for (int d = 0; d < 3; d++) {
            pos[d] = o->com[r][3 * m + d] + o->off[r][((size_t)m * o->natoms[r] + a) * 3 + d];
            th[d] = TWOPI * ((1.0 / o->L[d]) * pos[d]);
        }
        if (o->tri) { double s[3]; frac_tri(o, pos, s); for (int d = 0; d < 3; d++) th[d] = TWOPI * s[d]; }
        phase_factors_1d(&PH(o, px, 0, r, m, a, 0), th[0], o->kmax[0]);
        phase_factors_1d(&PH(o, py, 1, r, m, a, 0), th[1], o->kmax[1]);
        phase_factors_1d(&PH(o, pz, 2, r, m, a, 0), th[2], o->kmax[2]);
    }
}
/* ComputeAllFourierTerms :340-356 */
static void all_fourier_terms(orc_t* o) {
    for (int r = 0; r < o->nres; r++) for (int m = 0; m < o->count[r]; m++) orc_single_mol_fourier(o, r, m);
}
/* SaveSingleMolFourierTerms :134-183 (x saved for k>=0 only) */
void orc_save_fourier(orc_t* o, int r, int m) {
    for (int a = 0; a < o->natoms[r]; a++) {
        for (int k = 0; k <= o->kmax[0]; k++) PHOLD(o, px_old, 0, a, k) = PH(o, px, 0, r, m, a, k);
        for (int k = -o->kmax[1]; k <= o->kmax[1]; k++) PHOLD(o, py_old, 1, a, k) = PH(o, py, 1, r, m, a, k);
        for (int k = -o->kmax[2]; k <= o->kmax[2]; k++) PHOLD(o, pz_old, 2, a, k) = PH(o, pz, 2, r, m, a, k);
    }
    memcpy(o->A_old, o->A, sizeof(cplx) * o->nk);
}
/* RestoreSingleMolFourier :205-253 */
void orc_restore_fourier(orc_t* o, int r, int m) {
    for (int a = 0; a < o->natoms[r]; a++) {
        for (int k = 0; k <= o->kmax[0]; k++) PH(o, px, 0, r, m, a, k) = PHOLD(o, px_old, 0, a, k);
        for (int k = -o->kmax[1]; k <= o->kmax[1]; k++) PH(o, py, 1, r, m, a, k) = PHOLD(o, py_old, 1, a, k);
        for (int k = -o->kmax[2]; k <= o->kmax[2]; k++) PH(o, pz, 2, r, m, a, k) = PHOLD(o, pz_old, 2, a, k);
    }
    memcpy(o->A, o->A_old, sizeof(cplx) * o->nk);
}
/* ReplaceFourierTermsSingleMol :276-322 */
void orc_replace_fourier(orc_t* o, int r, int i1, int i2) {
    for (int a = 0; a < o->natoms[r]; a++) {
        for (int k = 0; k <= o->kmax[0]; k++) PH(o, px, 0, r, i1, a, k) = PH(o, px, 0, r, i2, a, k);
        for (int k = -o->kmax[1]; k <= o->kmax[1]; k++) PH(o, py, 1, r, i1, a, k) = PH(o, py, 1, r, i2, a, k);
        for (int k = -o->kmax[2]; k <= o->kmax[2]; k++) PH(o, pz, 2, r, i1, a, k) = PH(o, pz, 2, r, i2, a, k);
    }
}

/* ---------------- reciprocal energy, ewald_energy.f90 --------------------------------------------- */

/* ComputeRecipAmplitude :40-74 */
static cplx recip_amplitude(const orc_t* o, int kx, int ky, int kz) {
    cplx amp = {0.0, 0.0};
    for (int r = 0; r < o->nres; r++)
        for (int m = 0; m < o->count[r]; m++)
            for (int a = 0; a < o->natoms[r]; a++) {
                double q = o->charges[r * o->max_atoms + a];
                cplx ph = cmul(cmul(PH(o, px, 0, r, m, a, kx), PH(o, py, 1, r, m, a, ky)), PH(o, pz, 2, r, m, a, kz));
                amp.re += q * ph.re; amp.im += q * ph.im;
            }
    return amp;
}
/* ComputeReciprocalEnergy :105-148; A(k) is a local there (Q1); store_sk keeps it (init_sk=full) */
static double reciprocal_energy(orc_t* o, int store_sk) {
    double u = 0.0;
    for (int i = 0; i < o->nk; i++) {
        cplx amp = recip_amplitude(o, o->kx[i], o->ky[i], o->kz[i]);
        if (store_sk) o->A[i] = amp;
        double asq = amp.re * amp.re + amp.im * amp.im; /* amplitude_squared helper_utils.f90:125 */
        u += o->form[i] * o->W[i] * asq;
    }
    return u * EPS0_INV_eVA / KB_eVK * TWOPI / o->volume;
}

/* ComputeRecipEnergySingleMol :191-274.  Mutates A(k). */
double orc_recip_singlemol(orc_t* o, int r, int m, int creation_flag, int deletion_flag) {
    double u = 0.0;
    int na = o->natoms[r];
    const double* q = o->charges + r * o->max_atoms;
    for (int i = 0; i < o->nk; i++) {
        int kx = o->kx[i], ky = o->ky[i], kz = o->kz[i];
        cplx s = {0.0, 0.0};
        for (int a = 0; a < na; a++) {
            cplx pn = cmul(cmul(PH(o, px, 0, r, m, a, kx), PH(o, py, 1, r, m, a, ky)), PH(o, pz, 2, r, m, a, kz));
            cplx po = cmul(cmul(PHOLD(o, px_old, 0, a, kx), PHOLD(o, py_old, 1, a, ky)), PHOLD(o, pz_old, 2, a, kz));
            if (creation_flag) { s.re += q[a] * pn.re; s.im += q[a] * pn.im; }
            else if (deletion_flag) { s.re += q[a] * po.re; s.im += q[a] * po.im; }
            else { s.re += q[a] * (pn.re - po.re); s.im += q[a] * (pn.im - po.im); }
        }
        if (!creation_flag && deletion_flag) { o->A[i].re -= s.re; o->A[i].im -= s.im; }
        else { o->A[i].re += s.re; o->A[i].im += s.im; }
        double asq = o->A[i].re * o->A[i].re + o->A[i].im * o->A[i].im;
        u += o->form[i] * o->W[i] * asq;
    }
    return u * EPS0_INV_eVA / KB_eVK * TWOPI / o->volume;
}

/* ---------------- ComputeSystemEnergy energy_utils.f90:18-35 ---------------------------------------- */
void orc_system_energy(orc_t* o, double out[MMC_NE]) {
    pairwise_energy(o);
    /* ComputeEwaldSelf :307-330 (accumulates without zeroing, Q7; zeroed here so the call is idempotent) */
    o->energy[MMC_E_SELF] = 0.0;
    for (int r = 0; r < o->nres; r++) o->energy[MMC_E_SELF] += orc_self_singlemol(o, r) * o->count[r];
    /* ComputeEwaldRecip :270-287 */
    reciprocal_weights(o);
    all_fourier_terms(o);
    if (o->init_sk == MMC_INIT_SK_ZERO) memset(o->A, 0, sizeof(cplx) * o->nk);
    o->energy[MMC_E_RECIP] = reciprocal_energy(o, o->init_sk == MMC_INIT_SK_FULL);
    /* ComputeTotalIntraResidueCoulombEnergy :55-81 (active residues only) */
    o->energy[MMC_E_INTRA] = 0.0;
    for (int r = 0; r < o->nres; r++)
        if (o->is_active[r] == 1)
            for (int m = 0; m < o->count[r]; m++) o->energy[MMC_E_INTRA] += orc_intra_singlemol(o, r, m);
    o->energy[MMC_E_TOTAL] = o->energy[MMC_E_RECIP] + o->energy[MMC_E_NONCOUL] + o->energy[MMC_E_COUL] +
                             o->energy[MMC_E_SELF] + o->energy[MMC_E_INTRA];
    if (out) memcpy(out, o->energy, sizeof(double) * MMC_NE);
}

/* The set-up half of ComputeEwaldRecip (energy_utils.f90:270-287) WITHOUT the O(N K) amplitude sum and without the pair
 * sweep: W(k) and every molecule's 1-D phase tables (ComputeAllFourierTerms, ewald_phase.f90:340-356), tracked A(k) = 0,
 * tracked energies = 0.  With the reference's default (Q1: A(k) is never stored at start-up) the per-move records depend on
 * the start-up only through the tracked energies, so a chain prepared this way steps exactly like one whose tracked
 * energies were zeroed after ComputeSystemEnergy -- used where the start-up sums of a 50 000-atom framework (tens of
 * seconds of CPU) would dominate a bounded sample: bench.py's CPU leg and the bench-system parity runs. */
void orc_prepare_fourier(orc_t* o) {
    reciprocal_weights(o);
    all_fourier_terms(o);
    memset(o->A, 0, sizeof(cplx) * o->nk);
    memset(o->energy, 0, sizeof(double) * MMC_NE);
}

/* ---------------- Monte Carlo bookkeeping, monte_carlo_utils.f90 ------------------------------------ */

static void set_total(double e[MMC_NE], int all) {
    e[MMC_E_TOTAL] = all ? e[MMC_E_NONCOUL] + e[MMC_E_COUL] + e[MMC_E_RECIP] + e[MMC_E_SELF] + e[MMC_E_INTRA]
                         : e[MMC_E_NONCOUL] + e[MMC_E_COUL] + e[MMC_E_RECIP];
}
/* ComputeOldEnergy :347-396 */
void orc_compute_old_energy(orc_t* o, int r, int m, double old[MMC_NE], int creation, int deletion) {
    if (creation) {
        old[MMC_E_NONCOUL] = 0.0; old[MMC_E_COUL] = 0.0; old[MMC_E_SELF] = 0.0; old[MMC_E_INTRA] = 0.0;
        old[MMC_E_RECIP] = o->energy[MMC_E_RECIP];
        set_total(old, 1);
    } else if (deletion) {
        old[MMC_E_SELF] = orc_self_singlemol(o, r);
        old[MMC_E_INTRA] = orc_intra_singlemol(o, r, m);
        orc_pair_singlemol(o, r, m, &old[MMC_E_NONCOUL], &old[MMC_E_COUL]);
        old[MMC_E_RECIP] = o->energy[MMC_E_RECIP];
        set_total(old, 1);
    } else {
        old[MMC_E_SELF] = 0.0; old[MMC_E_INTRA] = 0.0; /* intent(out) fields left undefined by the reference */
        old[MMC_E_RECIP] = orc_recip_singlemol(o, r, m, 0, 0);
        orc_pair_singlemol(o, r, m, &old[MMC_E_NONCOUL], &old[MMC_E_COUL]);
        set_total(old, 0);
    }
}
/* ComputeNewEnergy :275-321.  Deletion passes is_creation = deletion_flag (:308, Q2); the
 * physical variant uses the is_deletion branch of ComputeRecipEnergySingleMol instead. */
void orc_compute_new_energy(orc_t* o, int r, int m, double nw[MMC_NE], int creation, int deletion) {
    if (creation) {
        orc_single_mol_fourier(o, r, m);
        nw[MMC_E_RECIP] = orc_recip_singlemol(o, r, m, 1, 0);
        orc_pair_singlemol(o, r, m, &nw[MMC_E_NONCOUL], &nw[MMC_E_COUL]);
        nw[MMC_E_SELF] = orc_self_singlemol(o, r);
        nw[MMC_E_INTRA] = orc_intra_singlemol(o, r, m);
        set_total(nw, 1);
    } else if (deletion) {
        nw[MMC_E_NONCOUL] = 0.0; nw[MMC_E_COUL] = 0.0; nw[MMC_E_SELF] = 0.0; nw[MMC_E_INTRA] = 0.0;
        if (o->deletion_mode == MMC_DELETION_REFERENCE) nw[MMC_E_RECIP] = orc_recip_singlemol(o, r, m, 1, 0);
        else nw[MMC_E_RECIP] = orc_recip_singlemol(o, r, m, 0, 1);
        set_total(nw, 1);
    } else {
        nw[MMC_E_SELF] = 0.0; nw[MMC_E_INTRA] = 0.0;
        orc_single_mol_fourier(o, r, m);
        nw[MMC_E_RECIP] = orc_recip_singlemol(o, r, m, 0, 0);
        orc_pair_singlemol(o, r, m, &nw[MMC_E_NONCOUL], &nw[MMC_E_COUL]);
        set_total(nw, 0);
    }
}
/* mc_acceptance_probability :184-221 (N read after the count update) */
static double acceptance_probability(const orc_t* o, const double old[MMC_NE], const double nw[MMC_NE], int r, int move) {
    double N = (double)(float)o->count[r]; /* real(primary%num_residues) is default (single) real */
    double V = o->volume, phi = o->fugacity[r], T = o->T;
    double de = nw[MMC_E_TOTAL] - old[MMC_E_TOTAL];
    switch (move) {
        case MMC_CREATION: return fmin(1.0, (phi * V / N) * exp(-de / T));
        case MMC_DELETION: return fmin(1.0, ((N + 1.0) / (phi * V)) * exp(-de / T));
        default: return fmin(1.0, exp(-de / T));
    }
}
/* AcceptMove :410-421 */
static void accept_move(orc_t* o, const double old[MMC_NE], const double nw[MMC_NE]) {
    o->energy[MMC_E_RECIP] = o->energy[MMC_E_RECIP] + nw[MMC_E_RECIP] - old[MMC_E_RECIP];
    o->energy[MMC_E_NONCOUL] = o->energy[MMC_E_NONCOUL] + nw[MMC_E_NONCOUL] - old[MMC_E_NONCOUL];
    o->energy[MMC_E_COUL] = o->energy[MMC_E_COUL] + nw[MMC_E_COUL] - old[MMC_E_COUL];
    o->energy[MMC_E_TOTAL] = o->energy[MMC_E_TOTAL] + nw[MMC_E_TOTAL] - old[MMC_E_TOTAL];
}
/* AcceptCreationMove create_molecule.f90:96-134 == AcceptDeletionMove delete_molecule.f90:126-170 (energies) */
static void accept_insdel(orc_t* o, const double old[MMC_NE], const double nw[MMC_NE]) {
    o->energy[MMC_E_RECIP] = nw[MMC_E_RECIP];
    o->energy[MMC_E_NONCOUL] = o->energy[MMC_E_NONCOUL] + nw[MMC_E_NONCOUL] - old[MMC_E_NONCOUL];
    o->energy[MMC_E_COUL] = o->energy[MMC_E_COUL] + nw[MMC_E_COUL] - old[MMC_E_COUL];
    o->energy[MMC_E_SELF] = o->energy[MMC_E_SELF] + nw[MMC_E_SELF] - old[MMC_E_SELF];
    o->energy[MMC_E_INTRA] = o->energy[MMC_E_INTRA] + nw[MMC_E_INTRA] - old[MMC_E_INTRA];
    o->energy[MMC_E_TOTAL] = o->energy[MMC_E_TOTAL] + nw[MMC_E_TOTAL] - old[MMC_E_TOTAL];
}

/* RotationMatrix helper_utils.f90:39-77 applied as in ApplyRandomRotation monte_carlo_utils.f90:30-66 */
static void apply_random_rotation(orc_t* o, int r, int m, int full) {
    int na = o->natoms[r];
    if (na == 1) return;
    double theta = full ? rand_uniform(o) * TWOPI : (rand_uniform(o) - 0.5) * o->rotation_step; /* ChooseRotationAngle :72-92 */
    int axis = (int)(rand_uniform(o) * 3.0) + 1;
    double c = cos(theta), s = sin(theta);
    double R[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    if (axis == 1) { R[1][1] = c; R[1][2] = -s; R[2][1] = s; R[2][2] = c; }
    else if (axis == 2) { R[0][0] = c; R[0][2] = s; R[2][0] = -s; R[2][2] = c; }
    else { R[0][0] = c; R[0][1] = -s; R[1][0] = s; R[1][1] = c; }
    for (int a = 0; a < na; a++) {
        double* v = o->off[r] + ((size_t)m * na + a) * 3;
        double w[3];
        for (int i = 0; i < 3; i++) w[i] = R[i][0] * v[0] + R[i][1] * v[1] + R[i][2] * v[2]; /* matmul */
        v[0] = w[0]; v[1] = w[1]; v[2] = w[2];
    }
}

static void fill_record(mmc_move_record* rec, int kind, int r, int m, int acc, const double* old, const double* nw, double p) {
    if (!rec) return;
    rec->kind = kind; rec->res = r; rec->mol = m; rec->accepted = acc; rec->prob = p;
    memcpy(rec->old_e, old, sizeof(double) * MMC_NE); memcpy(rec->new_e, nw, sizeof(double) * MMC_NE);
}

/* Translation translation.f90:36-112 */
static void translation(orc_t* o, int r, int m, mmc_move_record* rec) {
    if (m < 0) return; /* molecule_index == 0 */
    double old[MMC_NE], nw[MMC_NE], com_old[3];
    o->cnt.trial_translations++;
    orc_save_fourier(o, r, m); memcpy(com_old, o->com[r] + 3 * m, sizeof(com_old)); /* SaveMoleculeState :444 */
    orc_compute_old_energy(o, r, m, old, 0, 0);
    double u[3]; for (int d = 0; d < 3; d++) u[d] = rand_uniform(o) - 0.5; /* rand_symmetric(3) */
    for (int d = 0; d < 3; d++) o->com[r][3 * m + d] = o->com[r][3 * m + d] + u[d] * o->translation_step;
    apply_pbc(o, o->com[r] + 3 * m);
    orc_compute_new_energy(o, r, m, nw, 0, 0);
    double p = acceptance_probability(o, old, nw, r, MMC_TRANSLATION);
    int acc = rand_uniform(o) <= p;
    if (acc) { accept_move(o, old, nw); o->cnt.translations++; }
    else { memcpy(o->com[r] + 3 * m, com_old, sizeof(com_old)); orc_restore_fourier(o, r, m); } /* RejectMoleculeMove :485 */
    fill_record(rec, MMC_TRANSLATION, r, m, acc, old, nw, p);
}
/* Rotation rotation.f90:34-75 */
static void rotation(orc_t* o, int r, int m, mmc_move_record* rec) {
    int na = o->natoms[r];
    if (na == 1 || m < 0) return;
    double old[MMC_NE], nw[MMC_NE];
    o->cnt.trial_rotations++;
    orc_save_fourier(o, r, m); memcpy(o->off_old, o->off[r] + (size_t)m * na * 3, sizeof(double) * 3 * na);
    orc_compute_old_energy(o, r, m, old, 0, 0);
    apply_random_rotation(o, r, m, 0);
    orc_compute_new_energy(o, r, m, nw, 0, 0);
    double p = acceptance_probability(o, old, nw, r, MMC_ROTATION);
    int acc = rand_uniform(o) <= p;
    if (acc) { accept_move(o, old, nw); o->cnt.rotations++; }
    else { memcpy(o->off[r] + (size_t)m * na * 3, o->off_old, sizeof(double) * 3 * na); orc_restore_fourier(o, r, m); }
    fill_record(rec, MMC_ROTATION, r, m, acc, old, nw, p);
}
/* CreateMolecule create_molecule.f90:41-207 (no-reservoir branch: geometry of slot 1 + full rotation, Q9) */
static int create_molecule(orc_t* o, int r, int m, mmc_move_record* rec) {
    int na = o->natoms[r];
    if (m >= o->cap[r]) return 1; /* CheckMoleculeIndex output_utils.f90:604 -> AbortRun */
    double old[MMC_NE], nw[MMC_NE];
    o->cnt.trial_creations++;
    orc_compute_old_energy(o, r, m, old, 1, 0);
    o->count[r]++;
    orc_save_fourier(o, r, m);
    { double u[3], p[3];                                               /* :180-182: bounds(:,1) + matmul(matrix, u) */
      for (int d = 0; d < 3; d++) u[d] = rand_uniform(o);
      if (o->tri) { cart_tri(o, u, p); for (int d = 0; d < 3; d++) o->com[r][3 * m + d] = o->lo[d] + p[d]; }
      else for (int d = 0; d < 3; d++) o->com[r][3 * m + d] = o->lo[d] + o->L[d] * u[d]; }
    int rsv = -1;
    if (o->has_reservoir) { /* :185-193: geometry of a randomly drawn reservoir molecule, no rotation */
        rsv = (int)(rand_uniform(o) * o->rsv_count[r]);
        memcpy(o->off[r] + (size_t)m * na * 3, o->rsv_off[r] + (size_t)rsv * na * 3, sizeof(double) * 3 * na);
    } else {
        if (m != 0) memcpy(o->off[r] + (size_t)m * na * 3, o->off[r], sizeof(double) * 3 * na); /* :198-199 */
        apply_random_rotation(o, r, m, 1);
    }
    orc_compute_new_energy(o, r, m, nw, 1, 0);
    double p = acceptance_probability(o, old, nw, r, MMC_CREATION);
    int acc = rand_uniform(o) <= p;
    if (acc) {
        accept_insdel(o, old, nw); o->cnt.creations++;
        if (o->has_reservoir && o->rsv_count[r] > 0) { /* AcceptCreationMove :118-130: pool slot <- last pool slot */
            const int last = o->rsv_count[r] - 1;
            memmove(o->rsv_off[r] + (size_t)rsv * na * 3, o->rsv_off[r] + (size_t)last * na * 3, sizeof(double) * 3 * na);
            o->rsv_count[r]--;
        }
    }
    else { o->count[r]--; orc_restore_fourier(o, r, m); } /* RejectCreationMove :141 */
    fill_record(rec, MMC_CREATION, r, m, acc, old, nw, p);
    return 0;
}
/* DeleteMolecule delete_molecule.f90:41-199 */
static void delete_molecule(orc_t* o, int r, int m, mmc_move_record* rec) {
    int na = o->natoms[r];
    if (o->count[r] == 0) return;
    double old[MMC_NE], nw[MMC_NE];
    o->cnt.trial_deletions++;
    orc_compute_old_energy(o, r, m, old, 0, 1);
    orc_save_fourier(o, r, m);
    memcpy(o->com_old, o->com[r] + 3 * m, sizeof(double) * 3);
    memcpy(o->off_old, o->off[r] + (size_t)m * na * 3, sizeof(double) * 3 * na);
    int last = o->count[r] - 1;
    /* RemoveMolecule :99-118 */
    memmove(o->com[r] + 3 * m, o->com[r] + 3 * last, sizeof(double) * 3);
    memmove(o->off[r] + (size_t)m * na * 3, o->off[r] + (size_t)last * na * 3, sizeof(double) * 3 * na);
    orc_replace_fourier(o, r, m, last);
    o->count[r]--;
    orc_compute_new_energy(o, r, m, nw, 0, 1);
    double p = acceptance_probability(o, old, nw, r, MMC_DELETION);
    int acc = rand_uniform(o) <= p;
    if (acc) {
        accept_insdel(o, old, nw); o->cnt.deletions++;
        if (o->has_reservoir) { /* AcceptDeletionMove :148-166: three draws place the molecule in the reservoir box (its COM
                                  * is not used by the path); the offsets appended are those of slot last_molecule_index */
            rand_uniform(o); rand_uniform(o); rand_uniform(o);
            if (o->rsv_count[r] < o->rsv_cap[r]) {
                memcpy(o->rsv_off[r] + (size_t)o->rsv_count[r] * na * 3, o->off[r] + (size_t)last * na * 3, sizeof(double) * 3 * na);
                o->rsv_count[r]++;
            }
        }
    }
    else { /* RejectDeletionMove :178-199 */
        o->count[r]++;
        memcpy(o->com[r] + 3 * m, o->com_old, sizeof(double) * 3);
        memcpy(o->off[r] + (size_t)m * na * 3, o->off_old, sizeof(double) * 3 * na);
        orc_restore_fourier(o, r, m);
    }
    fill_record(rec, MMC_DELETION, r, m, acc, old, nw, p);
}

/* Swap (identity exchange), SURVEY 8(f).1.  The reference only ships the acceptance rule
 * (mc_acceptance_probability_swap, monte_carlo_utils.f90:228-268: min(1, phi_new/phi_old * N_old/(N_new+1) * exp(-dE/T)),
 * counts read before the move); the move itself is composed here from the reference's own deletion and creation
 * building blocks: ComputeOldEnergy(deletion) of (rA, m), physical removal (is_deletion branch of
 * ComputeRecipEnergySingleMol + RemoveMolecule), insertion of a type-rB molecule at the SAME centre of mass with the
 * slot-1 template geometry and a random orientation (create_molecule.f90:195-199), ComputeNewEnergy(creation).
 * Draws: type pick (only if more than two active types exist), rotation angle, axis, accept. */
static int swap_molecule(orc_t* o, int rA, int m, mmc_move_record* rec) {
    int n_active = 0;
    for (int i = 0; i < o->nres; i++) if (o->is_active[i] == 1) n_active++;
    if (n_active < 2 || m < 0) return 0;
    int pick = n_active > 2 ? (int)(rand_uniform(o) * (n_active - 1)) : 0;
    if (pick > n_active - 2) pick = n_active - 2;
    int rB = -1;
    for (int i = 0, k = 0; i < o->nres; i++) if (o->is_active[i] == 1 && i != rA) { if (k == pick) { rB = i; break; } k++; }
    const int mB = o->count[rB], naA = o->natoms[rA], naB = o->natoms[rB];
    if (mB >= o->cap[rB]) return 1;
    double old[MMC_NE], nw[MMC_NE];
    o->cnt.trial_swaps++;
    const double N_old = (double)o->count[rA], N_new = (double)o->count[rB];
    orc_compute_old_energy(o, rA, m, old, 0, 1);
    orc_save_fourier(o, rA, m);                       /* phases of (rA, m) and A(k) */
    memcpy(o->com_old, o->com[rA] + 3 * m, sizeof(double) * 3);
    memcpy(o->off_old, o->off[rA] + (size_t)m * naA * 3, sizeof(double) * 3 * naA);
    orc_recip_singlemol(o, rA, m, 0, 1);              /* A(k) -= sum q phi(old) */
    const int last = o->count[rA] - 1;
    memmove(o->com[rA] + 3 * m, o->com[rA] + 3 * last, sizeof(double) * 3);
    memmove(o->off[rA] + (size_t)m * naA * 3, o->off[rA] + (size_t)last * naA * 3, sizeof(double) * 3 * naA);
    orc_replace_fourier(o, rA, m, last);
    o->count[rA]--;
    /* the new molecule: same COM, template geometry of slot 1 of its type, random orientation */
    memcpy(o->com[rB] + 3 * mB, o->com_old, sizeof(double) * 3);
    if (mB != 0) memcpy(o->off[rB] + (size_t)mB * naB * 3, o->off[rB], sizeof(double) * 3 * naB);
    o->count[rB]++;
    apply_random_rotation(o, rB, mB, 1);
    orc_single_mol_fourier(o, rB, mB);
    nw[MMC_E_RECIP] = orc_recip_singlemol(o, rB, mB, 1, 0);
    orc_pair_singlemol(o, rB, mB, &nw[MMC_E_NONCOUL], &nw[MMC_E_COUL]);
    nw[MMC_E_SELF] = orc_self_singlemol(o, rB);
    nw[MMC_E_INTRA] = orc_intra_singlemol(o, rB, mB);
    set_total(nw, 1);
    const double de = nw[MMC_E_TOTAL] - old[MMC_E_TOTAL];
    const double p = fmin(1.0, (o->fugacity[rB] / o->fugacity[rA]) * (N_old / (N_new + 1.0)) * exp(-de / o->T));
    const int acc = rand_uniform(o) <= p;
    if (acc) { accept_insdel(o, old, nw); o->cnt.swaps++; }
    else {
        o->count[rB]--;
        o->count[rA]++;
        memcpy(o->com[rA] + 3 * m, o->com_old, sizeof(double) * 3);
        memcpy(o->off[rA] + (size_t)m * naA * 3, o->off_old, sizeof(double) * 3 * naA);
        orc_restore_fourier(o, rA, m);
    }
    fill_record(rec, MMC_SWAP, rA + 256 * rB, m, acc, old, nw, p);
    return 0;
}

/* AdjustMoveStepSizes monte_carlo_utils.f90:99-129 (single-precision ratios, Q6) */
void orc_adjust_step_sizes(orc_t* o) {
    if (!o->recalibrate) return;
    if (o->cnt.trial_translations > MIN_TRIALS_FOR_RECALIBRATION) {
        double acc = (double)((float)o->cnt.translations / (float)o->cnt.trial_translations);
        if (acc - TARGET_ACCEPTANCE > TOL_ACCEPTANCE) o->translation_step = fmin(o->translation_step * 1.05, MAX_TRANSLATION_STEP);
        else if (acc - TARGET_ACCEPTANCE < TOL_ACCEPTANCE) o->translation_step = fmax(o->translation_step * 0.95, MIN_TRANSLATION_STEP);
    }
    if (o->cnt.trial_rotations > MIN_TRIALS_FOR_RECALIBRATION) {
        double acc = (double)((float)o->cnt.rotations / (float)o->cnt.trial_rotations);
        if (acc - TARGET_ACCEPTANCE > TOL_ACCEPTANCE) o->rotation_step = fmin(o->rotation_step * 1.05, MAX_ROTATION_ANGLE);
        else if (acc - TARGET_ACCEPTANCE < TOL_ACCEPTANCE) o->rotation_step = fmin(o->rotation_step * 1.95, MIN_ROTATION_ANGLE);
    }
}

/* one iteration of the step loop of MonteCarloLoop, monte_carlo.f90:42-76 */
static int mc_step(orc_t* o, mmc_move_record* rec) {
    if (rec) memset(rec, 0, sizeof(*rec));
    /* PickRandomResidueType monte_carlo_utils.f90:136-164 */
    int n_active = 0, r = -1;
    for (int i = 0; i < o->nres; i++) if (o->is_active[i] == 1) n_active++;
    if (n_active == 0) return 2;
    int pick = (int)(rand_uniform(o) * n_active);
    for (int i = 0, k = 0; i < o->nres; i++) if (o->is_active[i] == 1) { if (k == pick) { r = i; break; } k++; }
    /* PickRandomMoleculeIndex :170-182 */
    int m = -1;
    if (o->count[r] != 0) { m = (int)(rand_uniform(o) * o->count[r]); if (m > o->count[r] - 1) m = o->count[r] - 1; }
    double draw = rand_uniform(o);
    if (draw <= o->p_translation) translation(o, r, m, rec);
    else if (draw <= o->p_rotation + o->p_translation) rotation(o, r, m, rec);
    else if (draw <= o->p_rotation + o->p_translation + o->p_swap) return swap_molecule(o, r, m, rec);
    else {
        if (rand_uniform(o) <= PROB_CREATE_DELETE) return create_molecule(o, r, o->count[r], rec);
        else delete_molecule(o, r, m, rec);
    }
    return 0;
}

/* Runs n_steps MC steps.  tape != NULL: uniforms read sequentially (strict mode);
 * else Philox (seed, chain, step0 + i, draw).  Returns 0, or 1 if capacity was exceeded. */
int orc_run(orc_t* o, int64_t n_steps, const double* tape, int64_t tape_len, uint64_t seed, uint32_t chain,
            uint64_t step0, mmc_move_record* recs, int64_t* tape_used) {
    o->rng.tape = tape; o->rng.pos = 0; o->rng.len = tape_len; o->rng.seed = seed; o->rng.chain = chain;
    int rc = 0;
    for (int64_t i = 0; i < n_steps; i++) {
        o->rng.step = step0 + (uint64_t)i; o->rng.draw = 0;
        rc = mc_step(o, recs ? recs + i : NULL);
        if (rc) break;
    }
    if (tape_used) *tape_used = o->rng.pos;
    return rc;
}

/* strict-mode single moves for kernel-level tests: returns records, uses the tape for draws */
int orc_single_move(orc_t* o, int kind, int r, int m, const double* tape, int64_t tape_len, mmc_move_record* rec) {
    o->rng.tape = tape; o->rng.pos = 0; o->rng.len = tape_len;
    if (rec) memset(rec, 0, sizeof(*rec));
    switch (kind) {
        case MMC_TRANSLATION: translation(o, r, m, rec); break;
        case MMC_ROTATION: rotation(o, r, m, rec); break;
        case MMC_CREATION: return create_molecule(o, r, o->count[r], rec);
        case MMC_DELETION: delete_molecule(o, r, m, rec); break;
        case MMC_SWAP: return swap_molecule(o, r, m, rec);
        default: return 3;
    }
    return 0;
}
