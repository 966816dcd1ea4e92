/* fortran_glue_probe.c -- TEST SUPPORT: a line-for-line C transcription of fortran/maniac_gpu_glue.f90 working on the
 * reference's REAL array shapes in column-major (Fortran) order, so that the C ABI is exercised exactly the way the
 * compile-untested Fortran glue would exercise it (neither the build image nor the GPU box has a Fortran compiler).
 *
 * Array shapes, as allocated by the reference (input_parser.f90:245-266, NB_MAX_MOLECULE = 5000, parameters.f90:8):
 *   mol_com(3, nres, NB_MAX)                 site_offset(3, nres, NB_MAX, max_atom)
 *   atom_charges(nres, max_atom)             atom_types(nres, max_atom)
 *   coeff%sigma / epsilon(nres, nres, max_atom, max_atom)          num_residues(nres)
 * Fortran element (i1, i2, ..) with 1-based indices lives at (i1-1) + n1*((i2-1) + n2*((i3-1) + ..)).
 * The functions take 1-based indices where the Fortran routines do.  The library is reached through function pointers
 * (the caller dlopens libmaniacgpu.so), so this file has no link-time dependency on CUDA.
 */
#include <stdlib.h>
#include <string.h>
#include "maniac_gpu.h"

typedef struct {
    /* module globals of simulation_state.f90 that the glue reads (column-major, caller-owned) */
    int nres, max_atom, ntypes, nb_max;              /* nb%type_residue, nb%max_atom_in_residue, primary%num_atomtypes, NB_MAX_MOLECULE */
    int* atom_in_residue;                            /* nb%atom_in_residue(nres) */
    int* is_active;                                  /* input%is_active(nres) */
    int* num_residues;                               /* primary%num_residues(nres) */
    double* fugacity;                                /* input%fugacity(nres) */
    double* atom_charges; int* atom_types;           /* (nres, max_atom) */
    double* sigma; double* epsilon;                  /* coeff%sigma/epsilon(nres, nres, max_atom, max_atom) */
    double* mol_com;                                 /* primary%mol_com(3, nres, NB_MAX) */
    double* site_offset;                             /* primary%site_offset(3, nres, NB_MAX, max_atom) */
    double bounds_lo[3], matrix_diag[3], tilt[3];
    double alpha, cutoff, temp_K, translation_step, rotation_step_angle, p_translation, p_rotation, p_swap;
    int kmax[3];
    double energy[6];                                /* module-level `energy` (energy_type) in the C ABI's order */
    /* packed copies (p_* of the glue) */
    int *p_natoms, *p_active, *p_count, *p_cap, *p_types;
    double *p_q, *p_sig, *p_eps, *p_com, *p_off, *p_fug;
    int n_com, n_off;
    /* the library */
    int (*init)(mmc_t**, const mmc_desc*);
    int (*system_energy)(mmc_t*, int, double*);
    int (*trial_energy)(mmc_t*, int, int, int, int, const double*, const double*, double*, double*);
    int (*trial_commit)(mmc_t*, int, int);
    int (*sync_to_host)(mmc_t*, int, int*, double*, double*, double*);
    void (*destroy)(mmc_t*);
    mmc_t* gpu;
} glue_state;

/* column-major addressing, 1-based like the Fortran source */
#define MOL_COM(g, d, r, m) (g)->mol_com[((d) - 1) + 3 * (((r) - 1) + (size_t)(g)->nres * ((m) - 1))]
#define SITE_OFF(g, d, r, m, a) (g)->site_offset[((d) - 1) + 3 * (((r) - 1) + (size_t)(g)->nres * (((m) - 1) + (size_t)(g)->nb_max * ((a) - 1)))]
#define RA(g, arr, r, a) (g)->arr[((r) - 1) + (size_t)(g)->nres * ((a) - 1)]
#define COEFF(g, arr, r, r2, a, a2) (g)->arr[((r) - 1) + (size_t)(g)->nres * (((r2) - 1) + (size_t)(g)->nres * (((a) - 1) + (size_t)(g)->max_atom * ((a2) - 1)))]

/* pack_coordinates (maniac_gpu_glue.f90:33-53) */
void glue_pack_coordinates(glue_state* g) {
    int r, m, a, im, ia, nmol = 0, natom = 0;
    for (r = 1; r <= g->nres; r++) {
        nmol += g->num_residues[r - 1];
        natom += g->num_residues[r - 1] * g->atom_in_residue[r - 1];
    }
    free(g->p_com); free(g->p_off);
    g->p_com = (double*)calloc(3 * (size_t)(nmol > 1 ? nmol : 1), sizeof(double));
    g->p_off = (double*)calloc(3 * (size_t)(natom > 1 ? natom : 1), sizeof(double));
    g->n_com = nmol; g->n_off = natom;
    im = 0; ia = 0;
    for (r = 1; r <= g->nres; r++)
        for (m = 1; m <= g->num_residues[r - 1]; m++) {
            for (int d = 1; d <= 3; d++) g->p_com[3 * im + d - 1] = MOL_COM(g, d, r, m);
            im++;
            for (a = 1; a <= g->atom_in_residue[r - 1]; a++) {
                for (int d = 1; d <= 3; d++) g->p_off[3 * ia + d - 1] = SITE_OFF(g, d, r, m, a);
                ia++;
            }
        }
}

/* the packing half of gpu_attach (maniac_gpu_glue.f90:55-88) */
void glue_pack_tables(glue_state* g) {
    const int nres = g->nres, maxat = g->max_atom, ntypes = g->ntypes;
    int r, r2, a, a2, ti, tj;
    g->p_natoms = (int*)calloc(nres, sizeof(int)); g->p_active = (int*)calloc(nres, sizeof(int));
    g->p_count = (int*)calloc(nres, sizeof(int)); g->p_cap = (int*)calloc(nres, sizeof(int));
    g->p_fug = (double*)calloc(nres, sizeof(double));
    g->p_q = (double*)calloc((size_t)nres * maxat, sizeof(double)); g->p_types = (int*)calloc((size_t)nres * maxat, sizeof(int));
    g->p_sig = (double*)calloc((size_t)ntypes * ntypes, sizeof(double)); g->p_eps = (double*)calloc((size_t)ntypes * ntypes, sizeof(double));
    for (size_t i = 0; i < (size_t)nres * maxat; i++) g->p_types[i] = 1;
    for (r = 1; r <= nres; r++) {
        g->p_natoms[r - 1] = g->atom_in_residue[r - 1];
        g->p_active[r - 1] = g->is_active[r - 1] == 1 ? 1 : 0;
        g->p_count[r - 1] = g->num_residues[r - 1];
        g->p_cap[r - 1] = g->is_active[r - 1] == 1 ? g->nb_max : g->num_residues[r - 1];
        g->p_fug[r - 1] = g->fugacity[r - 1];
        for (a = 1; a <= g->atom_in_residue[r - 1]; a++) {
            g->p_q[(size_t)(r - 1) * maxat + a - 1] = RA(g, atom_charges, r, a);
            g->p_types[(size_t)(r - 1) * maxat + a - 1] = RA(g, atom_types, r, a);
        }
    }
    for (r = 1; r <= nres; r++)
        for (r2 = 1; r2 <= nres; r2++)
            for (a = 1; a <= g->atom_in_residue[r - 1]; a++)
                for (a2 = 1; a2 <= g->atom_in_residue[r2 - 1]; a2++) {
                    ti = RA(g, atom_types, r, a); tj = RA(g, atom_types, r2, a2);
                    g->p_sig[(size_t)(ti - 1) * ntypes + tj - 1] = COEFF(g, sigma, r, r2, a, a2);
                    g->p_eps[(size_t)(ti - 1) * ntypes + tj - 1] = COEFF(g, epsilon, r, r2, a, a2);
                }
    glue_pack_coordinates(g);
}

/* gpu_attach (maniac_gpu_glue.f90:55-110) */
int glue_attach(glue_state* g, int n_chains, int device, int init_sk, int deletion_mode) {
    mmc_desc d;
    memset(&d, 0, sizeof d);
    glue_pack_tables(g);
    d.nres = g->nres; d.max_atoms = g->max_atom; d.ntypes = g->ntypes; d.n_chains = n_chains;
    d.natoms = g->p_natoms; d.is_active = g->p_active; d.count = g->p_count; d.capacity = g->p_cap;
    d.charges = g->p_q; d.types = g->p_types; d.sigma = g->p_sig; d.epsilon = g->p_eps;
    d.com = g->p_com; d.offsets = g->p_off; d.fugacity = g->p_fug;
    for (int i = 0; i < 3; i++) { d.box_lo[i] = g->bounds_lo[i]; d.box_len[i] = g->matrix_diag[i]; d.kmax[i] = g->kmax[i]; d.tilt[i] = g->tilt[i]; }
    d.alpha = g->alpha; d.cutoff = g->cutoff; d.temperature = g->temp_K;
    d.translation_step = g->translation_step; d.rotation_step = g->rotation_step_angle;
    d.p_translation = g->p_translation; d.p_rotation = g->p_rotation; d.p_swap = g->p_swap;
    d.init_sk = init_sk; d.deletion_mode = deletion_mode; d.use_table = 0;   /* the glue passes 0, 0 (quirks Q1, Q2) */
    d.device = device; d.coulomb_xcut = 0.0;
    d.shard_rank = 0; d.shard_count = 1;
    d.reservoir_count = 0; d.reservoir_offsets = 0;
    return g->init(&g->gpu, &d);
}

/* gpu_system_energy */
int glue_system_energy(glue_state* g) { return g->system_energy(g->gpu, 0, g->energy); }

/* gpu_trial (maniac_gpu_glue.f90:128-141): the proposal is already in primary%mol_com / site_offset(:, res, mol, :) */
int glue_trial(glue_state* g, int kind, int res, int mol, double old_e[6], double new_e[6]) {
    double com[3];
    double* off = (double*)calloc(3 * (size_t)g->max_atom, sizeof(double));
    for (int d = 1; d <= 3; d++) com[d - 1] = MOL_COM(g, d, res, mol);
    for (int a = 1; a <= g->atom_in_residue[res - 1]; a++)
        for (int d = 1; d <= 3; d++) off[3 * a - 3 + d - 1] = SITE_OFF(g, d, res, mol, a);
    const int rc = g->trial_energy(g->gpu, 0, kind, res - 1, mol - 1, com, off, old_e, new_e);
    free(off);
    return rc;
}

int glue_commit(glue_state* g, int accepted) { return g->trial_commit(g->gpu, 0, accepted ? 1 : 0); }

/* gpu_sync_primary (maniac_gpu_glue.f90:163-189) */
int glue_sync_primary(glue_state* g, int chain) {
    int r, m, a, im, ia, ncap = 0, acap = 0;
    int* cnt = (int*)calloc(g->nres, sizeof(int));
    for (r = 1; r <= g->nres; r++) {
        const int cap = g->is_active[r - 1] == 1 ? g->nb_max : g->num_residues[r - 1];
        ncap += cap; acap += cap * g->atom_in_residue[r - 1];
    }
    double* com = (double*)calloc(3 * (size_t)ncap + 3, sizeof(double));
    double* off = (double*)calloc(3 * (size_t)acap + 3, sizeof(double));
    double e[6];
    const int rc = g->sync_to_host(g->gpu, chain, cnt, com, off, e);
    if (rc == 0) {
        im = 0; ia = 0;
        for (r = 1; r <= g->nres; r++) {
            if (g->is_active[r - 1] == 1) g->num_residues[r - 1] = cnt[r - 1];
            for (m = 1; m <= cnt[r - 1]; m++) {
                if (g->is_active[r - 1] == 1) for (int d = 1; d <= 3; d++) MOL_COM(g, d, r, m) = com[3 * im + d - 1];
                im++;
                for (a = 1; a <= g->atom_in_residue[r - 1]; a++) {
                    if (g->is_active[r - 1] == 1) for (int d = 1; d <= 3; d++) SITE_OFF(g, d, r, m, a) = off[3 * ia + d - 1];
                    ia++;
                }
            }
        }
        memcpy(g->energy, e, sizeof e);
    }
    free(cnt); free(com); free(off);
    return rc;
}

void glue_detach(glue_state* g) {
    if (g->gpu) g->destroy(g->gpu);
    g->gpu = 0;
    free(g->p_natoms); free(g->p_active); free(g->p_count); free(g->p_cap); free(g->p_types);
    free(g->p_q); free(g->p_sig); free(g->p_eps); free(g->p_com); free(g->p_off); free(g->p_fug);
    g->p_natoms = g->p_active = g->p_count = g->p_cap = g->p_types = 0;
    g->p_q = g->p_sig = g->p_eps = g->p_com = g->p_off = g->p_fug = 0;
}

size_t glue_sizeof_state(void) { return sizeof(glue_state); }
