/*
 * maniac_gpu.h -- C ABI of libmaniacgpu.so, the B200 (sm_100a) replacement for MANIAC's
 * per-move energy-difference path (reference: maniac-mc/maniac-mc @ 961f5c9).
 *
 * The reference has no FFI boundary; the path sits behind Fortran module procedures working
 * on module globals.  This header defines the boundary at the edge between the Monte Carlo
 * move code (translation.f90, rotation.f90, create_molecule.f90, delete_molecule.f90,
 * monte_carlo_utils.f90) and the energy routines (energy_utils.f90, ewald_phase.f90,
 * ewald_energy.f90, ewald_kvectors.f90, tabulated_utils.f90).  Every entry point names the
 * reference routine(s) it replaces.  The Fortran side binds these through ISO_C_BINDING
 * (fortran/maniac_gpu_iface.f90).
 *
 * Conventions
 *  - plain C types only; the caller owns its buffers, the library copies at the call.
 *  - residue (molecule type) and molecule indices are 0-based here; the Fortran wrapper
 *    subtracts 1.
 *  - every function returns 0 on success, non-zero on error; mmc_last_error() gives the text
 *    (the Fortran wrapper maps it to AbortRun, output_utils.f90:535).
 *  - energies are in Kelvin exactly as in the reference (energy_utils.f90:440, ewald_energy.f90:272).
 *  - an energy vector is double[6] = {non_coulomb, coulomb, recip_coulomb, ewald_self,
 *    intra_coulomb, total}  (type energy_state, simulation_state.f90:74-82).
 *  - one handle may hold many independent chains (pressure points x replicas) that share the
 *    framework, the LJ table and the k-vector set; chain 0 is "the" system in strict mode.
 *  - calls on one handle must be serialised by the caller (the reference is single threaded).
 *  - there is NO CPU fallback: without a CUDA device mmc_init fails.
 */
#ifndef MANIAC_GPU_H
#define MANIAC_GPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MMC_MAX_MOL_ATOMS 16 /* max atoms of a movable (active) molecule */
#define MMC_NE 6             /* components of an energy vector */

/* move types, parameters.f90:35-38 */
enum { MMC_CREATION = 1, MMC_DELETION = 2, MMC_TRANSLATION = 3, MMC_ROTATION = 4,
       MMC_SWAP = 5 /* identity exchange: advertised (README.md:22), parsed (input_parser.f90:446-452) and given an
                     * acceptance rule (monte_carlo_utils.f90:228-268) by the reference but never called
                     * (monte_carlo.f90:50-75); built here as SURVEY.md 8(f).1, off unless p_swap > 0 */ };

/* quirk switches: defaults reproduce the reference */
enum { MMC_INIT_SK_ZERO = 0, MMC_INIT_SK_FULL = 1 };           /* Q1 */
enum { MMC_DELETION_REFERENCE = 0, MMC_DELETION_PHYSICAL = 1 }; /* Q2 */

enum { MMC_E_NONCOUL = 0, MMC_E_COUL = 1, MMC_E_RECIP = 2, MMC_E_SELF = 3, MMC_E_INTRA = 4, MMC_E_TOTAL = 5 };

/* System description handed over once after PrecomputeTable (main.f90:26).
 * Packed layouts (the Fortran shim packs them from the strided simulation_state arrays):
 *   charges[r*max_atoms + a], types[r*max_atoms + a]   <- primary%atom_charges/atom_types(r,a)
 *   sigma/epsilon[(ti-1)*ntypes + (tj-1)]              <- coeff%sigma/epsilon compacted by atom type
 *   com[3*M], offsets[3*A]: molecules in (residue, molecule) order, atoms in (residue, molecule,
 *   atom) order                                        <- primary%mol_com / primary%site_offset
 */
typedef struct mmc_desc {
    int nres;              /* nb%type_residue */
    int max_atoms;         /* leading dimension of charges/types (>= max natoms) */
    int ntypes;            /* primary%num_atomtypes */
    int n_chains;          /* independent chains sharing this description (>=1) */
    const int* natoms;     /* [nres] nb%atom_in_residue */
    const int* is_active;  /* [nres] input%is_active */
    const int* count;      /* [nres] primary%num_residues */
    const int* capacity;   /* [nres] slots per active type (<= NB_MAX_MOLECULE, parameters.f90:8) */
    const double* charges; /* [nres*max_atoms] */
    const int* types;      /* [nres*max_atoms] 1-based atom types */
    const double* sigma;   /* [ntypes*ntypes] Angstrom */
    const double* epsilon; /* [ntypes*ntypes] Kelvin (parameters_parser.f90:82) */
    const double* com;     /* [3 * sum(count)] */
    const double* offsets; /* [3 * sum(count*natoms)] */
    double box_lo[3];      /* primary%bounds(:,1) */
    double box_len[3];     /* primary%matrix(d,d): cell edge lengths lx, ly, lz (box%type 1,2; type 3 adds `tilt` below) */
    double alpha;          /* ewald%alpha */
    double cutoff;         /* input%real_space_cutoff */
    double temperature;    /* input%temp_K */
    const double* fugacity;/* [nres] already converted to 1/A^3 (prepare_utils.f90:68) */
    int kmax[3];           /* ewald%kmax */
    double translation_step; /* input%translation_step */
    double rotation_step;    /* input%rotation_step_angle */
    double p_translation;    /* proba%translation */
    double p_rotation;       /* proba%rotation */
    int init_sk;           /* MMC_INIT_SK_* */
    int deletion_mode;     /* MMC_DELETION_* */
    int use_table;         /* 0: direct erfc/LJ (reference behaviour, parameters.f90:42); 1: tabulated_utils.f90 tables */
    int device;            /* CUDA device ordinal */
    double coulomb_xcut;   /* real-space Coulomb terms with alpha*r >= xcut are skipped by the per-move sweep:
                            * erfc(xcut) is below one ulp of the sums the reference forms over all minimum-image
                            * pairs (energy_utils.f90:425-433).  0 = default 6.0 (erfc = 2.2e-17, below the
                            * 1.1e-16 rounding of a unit term), max 7;
                            * negative = evaluate every pair literally as the reference does */
    int shard_rank;        /* k-sharded mode (ONE large system over several GPUs, one handle per GPU): this handle */
    int shard_count;       /* sweeps part shard_rank of shard_count of the tiles and of the k rows; 0 or 1 = off    */
    double p_swap;         /* proba%swap (input_parser.f90:446-452); 0 = the reference's de-facto behaviour */
    /* Reservoir box (has_reservoir, data_parser.f90:60-62): a pool of molecule geometries per active type.  With a
     * reservoir an insertion copies the site offsets of a randomly drawn pool molecule instead of rotating the slot-1
     * template (create_molecule.f90:185-193), an accepted insertion removes that molecule from the pool (:118-130) and
     * an accepted deletion appends one (delete_molecule.f90:148-166).  Batched mode keeps one pool per chain on the
     * device; in strict mode the host owns the pool and these stay NULL.  NULL = no reservoir. */
    const int* reservoir_count;      /* [nres] reservoir%num_residues */
    const double* reservoir_offsets; /* [3 * sum(reservoir_count*natoms)] reservoir%site_offset, packed like offsets */
    /* Triclinic cell (box%type 3): LAMMPS tilt factors xy, xz, yz (box%tilt, readers_utils.f90:238-240) of the cell
     * a = (lx, 0, 0), b = (xy, ly, 0), c = (xz, yz, lz) with lx, ly, lz = box_len; all zero = orthorhombic.  The
     * reference mixes row and column conventions for this cell (ApplyPBC does not even return a point
     * it need not wrap), so the library implements ONE consistent convention instead: fractional coordinates
     * s = (a b c)^-1 (r - box_lo), k = 2 pi (a b c)^-T n, nearest lattice image (ComputeDistance's 27-image search,
     * geometry_utils.f90:397-411, with a, b, c as the shifts), insertion at box_lo + (a b c) u. */
    double tilt[3];
} mmc_desc;

/* One record per trial move of a batched run (optional output). */
typedef struct mmc_move_record {
    int32_t kind;     /* MMC_* move type, 0 = no move happened (e.g. translation with N==0) */
    int32_t res;      /* residue type; swap: removed type + 256 * inserted type */
    int32_t mol;      /* molecule slot */
    int32_t accepted; /* 1/0 */
    double old_e[MMC_NE];
    double new_e[MMC_NE];
    double prob;      /* acceptance probability */
} mmc_move_record;

/* counter_type, simulation_state.f90:19-31 */
typedef struct mmc_counters {
    int64_t translations, rotations, creations, deletions;
    int64_t trial_translations, trial_rotations, trial_creations, trial_deletions;
    int64_t swaps, trial_swaps;   /* counter%swap, counter%trial_swap (simulation_state.f90:24,29) */
} mmc_counters;

typedef struct mmc_handle mmc_t;

/* ---- set-up ------------------------------------------------------------------------------ */

/* Ewald parameters from cut-off and tolerance: replaces ComputeEwaldParameters and
 * ComputeFourierIndices (prepare_utils.f90:169-214).  Host-only arithmetic (runs once). */
int mmc_ewald_setup(double cutoff, double tolerance, const double box_len[3],
                    double* alpha, int kmax[3], int* num_kvectors);

/* Build the device-resident state: SoA coordinates/charges/types, LJ table, k-vector set
 * (PrecomputeValidReciprocalVectors, ewald_kvectors.f90:44), weights W(k)
 * (ComputeReciprocalWeights, ewald_kvectors.f90:225), optional tables (PrecomputeTable,
 * tabulated_utils.f90:10).  Replaces AllocateArray (prepare_utils.f90:220). */
int mmc_init(mmc_t** handle, const mmc_desc* desc);
void mmc_destroy(mmc_t* handle);
const char* mmc_last_error(void);

int mmc_num_kvectors(const mmc_t* handle);
/* copies of the k-set: kx,ky,kz,form_factor (ewald%kvectors, ewald%form_factor) and W(k)
 * (ewald%recip_constants); any pointer may be NULL */
int mmc_get_kvectors(mmc_t* handle, int* kx, int* ky, int* kz, double* weight, double* form);

/* per-chain thermodynamic state (isotherm sweeps): fugacity[nres] in 1/A^3, temperature in K */
int mmc_set_chain_params(mmc_t* handle, int chain, const double* fugacity, double temperature);
/* overwrite / read back one chain's guests: count[nres], packed com/offsets as in mmc_desc */
int mmc_set_chain_state(mmc_t* handle, int chain, const int* count, const double* com, const double* offsets);
int mmc_get_chain_state(mmc_t* handle, int chain, int* count, double* com, double* offsets);
/* step sizes (input%translation_step / rotation_step_angle; AdjustMoveStepSizes stays on the host) */
/* Philox stream of a chain in batched mode: draws are keyed by (seed; step, chain id, draw).  Default id = chain
 * index.  Give every chain of a multi-GPU sweep its GLOBAL id (e.g. 1000 * point + replica) and its
 * trajectory no longer depends on which GPU or which slot it was dealt to. */
int mmc_set_chain_id(mmc_t* handle, int chain, uint32_t id);
int mmc_set_step_sizes(mmc_t* handle, int chain, double translation_step, double rotation_step);
/* AdjustMoveStepSizes (monte_carlo_utils.f90:99-129) for every chain from its own counters, enqueued on the library's
 * stream; call once per block when input%recalibrate_moves is set (monte_carlo.f90:78).  Literal, quirks included. */
int mmc_adjust_step_sizes(mmc_t* handle);
int mmc_get_step_sizes(mmc_t* handle, int chain, double* translation_step, double* rotation_step);

/* ---- single-point energy ----------------------------------------------------------------- */

/* ComputeSystemEnergy (energy_utils.f90:18): pair sweep i<j with intra-molecular exclusion,
 * Ewald self, full S(k) + reciprocal energy, intra-residue correction of active residues.
 * Stores the result as the chain's tracked `energy` and, with MMC_INIT_SK_FULL, the tracked
 * A(k) := S(k) (with MMC_INIT_SK_ZERO it is zeroed, the reference's de-facto behaviour). */
int mmc_system_energy(mmc_t* handle, int chain, double out[MMC_NE]);
/* The same for EVERY chain of the handle in one launch set with one synchronisation.  The inactive residues' share --
 * their structure factor (ComputeRecipAmplitude over the framework, ewald_energy.f90:40-74), their mutual pair energy and
 * Ewald self term (energy_utils.f90:270-330) -- is identical for all chains and is computed once per handle; each chain adds
 * its guests (A_chain(k) = S_framework(k) + S_guests(k)).  out: [n_chains][MMC_NE] or NULL. */
int mmc_system_energy_all(mmc_t* handle, double* out);

/* ---- strict mode: one trial move per call, host draws the random numbers -------------------- */

/* SaveMoleculeState + ComputeOldEnergy + <proposal applied by the host> + ComputeNewEnergy
 * (monte_carlo_utils.f90:444,347,275) in one fused device pass.
 *   kind = MMC_TRANSLATION / MMC_ROTATION: mol is the moved slot, com_new/offs_new its proposed
 *          geometry (3 and 3*natoms doubles).
 *   kind = MMC_CREATION: mol must equal the current count (new last slot).
 *   kind = MMC_DELETION: com_new/offs_new ignored (may be NULL).
 * Nothing is committed; old_e/new_e receive type(energy_state) old and new. */
int mmc_trial_energy(mmc_t* handle, int chain, int kind, int res, int mol,
                     const double* com_new, const double* offs_new,
                     double old_e[MMC_NE], double new_e[MMC_NE]);

/* AcceptMove / AcceptCreationMove / AcceptDeletionMove (accept=1) or RejectMoleculeMove /
 * RejectCreationMove / RejectDeletionMove (accept=0) for the pending trial
 * (monte_carlo_utils.f90:410,485; create_molecule.f90:96,141; delete_molecule.f90:126,178).
 * A(k) is double buffered: accept flips the buffer, reject costs nothing
 * (replaces Save/RestoreSingleMolFourier, ewald_phase.f90:134,205). */
int mmc_trial_commit(mmc_t* handle, int chain, int accept);
/* Strict-mode swap trial (MMC_SWAP): molecule (res_old, mol) is removed and a molecule of type res_new with the
 * proposed geometry (com_new, offs_new: 3 and 3*natoms(res_new) doubles) is created in slot count(res_new).
 * old_e / new_e as for a deletion / a creation; commit with mmc_trial_commit. */
int mmc_trial_energy_swap(mmc_t* handle, int chain, int res_old, int mol, int res_new,
                          const double* com_new, const double* offs_new, double old_e[MMC_NE], double new_e[MMC_NE]);

/* One strict-mode move in ONE launch: mmc_trial_energy + the accept test + mmc_trial_commit.  `u` is the uniform the move
 * routine draws for its Metropolis test (`if (rand_uniform() <= probability)`, translation.f90:66, rotation.f90:62,
 * create_molecule.f90:77, delete_molecule.f90:80): no draw happens between the proposal and that one, so drawing it before
 * the call leaves the random stream untouched.  The library evaluates mc_acceptance_probability (monte_carlo_utils.f90:184-221),
 * takes the decision u <= prob and applies Accept* / Reject*.  res_new is read for kind = MMC_SWAP only.
 * Outputs: old_e / new_e (type(energy_state)), *prob, *accepted = 1 / 0; any pointer may be NULL. */
int mmc_trial_step(mmc_t* handle, int chain, int kind, int res, int mol, int res_new,
                   const double* com_new, const double* offs_new, double u,
                   double old_e[MMC_NE], double new_e[MMC_NE], double* prob, int* accepted);

/* once per block (PrintStatus / UpdateFiles, monte_carlo.f90:82-84): tracked energies,
 * counts and packed coordinates of one chain */
int mmc_sync_to_host(mmc_t* handle, int chain, int* count, double* com, double* offsets,
                     double energies[MMC_NE]);
/* Bulk form of the per-block exchange for many chains (isotherm sweeps): the host keeps, per chain,
 * fixed-capacity arrays like the reference's primary%mol_com(3,nres,NB_MAX) / site_offset
 * (input_parser.f90:250-251), restricted to the ACTIVE residues:
 *   count  [n_chains][nres]
 *   com    [n_chains][g_mols ][3]   slot of (active residue r, molecule m) = mol_base[r] + m
 *   offsets[n_chains][g_atoms][3]   slot of (r, m, a) = atom_base[r] + m*natoms[r] + a
 * with g_mols = sum_r capacity[r], g_atoms = sum_r capacity[r]*natoms[r] over active residues in
 * residue order (mmc_chain_capacity returns them).  Host buffers may be pinned; copies are
 * asynchronous on the handle's stream, the calls return after the stream has drained.
 * mmc_download_chains also returns tracked energies [n_chains][MMC_NE] and counters [n_chains]
 * (any pointer may be NULL). */
int mmc_chain_capacity(const mmc_t* handle, int* g_mols, int* g_atoms);
int mmc_upload_chains(mmc_t* handle, const int* count, const double* com, const double* offsets);
int mmc_download_chains(mmc_t* handle, int* count, double* com, double* offsets, double* energies,
                        mmc_counters* counters);
/* Per-block output without stalling the chains (UpdateFiles / PrintStatus, monte_carlo.f90:82-84, run between blocks;
 * the reference's writers are write_utils.f90:13-188).  mmc_snapshot_begin queues, behind the work already on the
 * handle's stream, a device-side copy of every chain's counts, coordinates, energies and counters, and copies it to
 * library-owned pinned host memory on a second stream: the next mmc_run_batch_async may be issued immediately.
 * slot = 0 or 1 (two snapshots may be in flight).  mmc_snapshot_wait blocks until THAT copy has landed (not until the
 * handle's stream drains) and returns pointers laid out like mmc_download_chains' arguments (count [n_chains][nres]);
 * they stay valid until the next mmc_snapshot_begin on the same slot or mmc_destroy. */
int mmc_snapshot_begin(mmc_t* handle, int slot);
int mmc_snapshot_wait(mmc_t* handle, int slot, const int** count, const double** com, const double** offsets,
                      const double** energies, const mmc_counters** counters);
/* tracked A(k) as (re,im) pairs, 2*K doubles (ewald%recip_amplitude) */
/* reservoir pool of one chain (batched mode): count[nres], offsets packed like mmc_desc.reservoir_offsets
 * (capacity reservoir_count + capacity molecules per type) */
int mmc_get_reservoir(mmc_t* handle, int chain, int* count, double* offsets);
int mmc_get_sk(mmc_t* handle, int chain, double* sk);

/* ---- batched mode: the inner `do current_step` loop of MonteCarloLoop on the device ------- */

/* Runs n_steps Monte Carlo steps (monte_carlo.f90:42-76) on EVERY chain: one CTA (or one thread-block cluster) owns a
 * chain and loops over the steps inside the kernel, 128 steps per launch, the launches replayed from a CUDA graph.
 * Random numbers: if tape != NULL chain c reads its uniforms sequentially from
 * tape[c*tape_stride ...] (strict replay of a recorded stream, draw order of SURVEY 3.8);
 * otherwise Philox4x32-10 keyed by seed with counter (step, chain, draw).
 * records (optional, may be NULL): n_chains*n_steps entries, chain-major.
 * counters (optional): n_chains entries, accumulated since mmc_init.
 * One deliberate difference from ComputeOldEnergy (monte_carlo_utils.f90:347-396), which re-sums the selected molecule's
 * pair energy with every other atom for the old geometry: the framework never moves, so these kernels keep, per guest
 * molecule, the framework part of that sum from the move that put the molecule where it is and re-sum only the guest-guest
 * part; entries they do not know (start-up, strict-mode accepts, mmc_set_chain_state, changed coordinates in
 * mmc_upload_chains) are re-summed.  Same energies, probabilities and accept sequences to 1e-10; MANIAC_FW_CACHE=0 in the
 * environment at mmc_init restores the re-summation. */
int mmc_run_batch(mmc_t* handle, int64_t n_steps, uint64_t seed,
                  const double* tape, int64_t tape_stride,
                  mmc_move_record* records, mmc_counters* counters);
/* same, but only enqueues on the handle's stream (device-resident benchmarking);
 * pair with mmc_synchronize. records are left on the device. */
int mmc_run_batch_async(mmc_t* handle, int64_t n_steps, uint64_t seed);
int mmc_synchronize(mmc_t* handle);
/* number of uniforms chain consumed from its tape in the last mmc_run_batch */
int64_t mmc_tape_used(mmc_t* handle, int chain);

/* ---- k-sharded mode: one very large system (or a few chains of it) spread over G GPUs ----------------------------
 * (no reference counterpart: the reference is a single-threaded CPU code; SURVEY.md 8e "optional single-huge-system
 * mode").  Every rank creates a handle from the SAME description with shard_rank = r, shard_count = G: it holds all
 * coordinates, sweeps 1/G of the framework tiles and owns 1/G of the (kx, ky) rows of A(k).  One MC step of all
 * chains = begin (proposal from the shared Philox counters + partial sweeps, six partial sums per chain left in the
 * caller's DEVICE buffer, all enqueued on mmc_stream) -> caller all-reduces dev_sums in place (ncclAllReduce, sum,
 * double, 6 x n_chains) on that stream -> end (energy assembly, accept draw, commit: identical on every rank, so no
 * broadcast follows).  records (host, [n_chains], may be NULL) receives the step's move records. */
int mmc_shard_step_begin(mmc_t* handle, uint64_t seed, double* dev_sums);
int mmc_shard_step_end(mmc_t* handle, const double* dev_sums, mmc_move_record* records);
/* strict-mode variant (host proposal, host decision): mmc_trial_energy split around the all-reduce of dev_sums[6];
 * follow with mmc_trial_commit on every rank. */
int mmc_shard_trial_begin(mmc_t* handle, int chain, int kind, int res, int mol,
                          const double* com_new, const double* offs_new, double* dev_sums);
int mmc_shard_trial_end(mmc_t* handle, int chain, const double* dev_sums, double old_e[MMC_NE], double new_e[MMC_NE]);
/* k-vectors whose A(k) this handle owns and keeps current: [k_first, k_first + k_count) of the k list */
int mmc_shard_range(const mmc_t* handle, int* k_first, int* k_count);

/* ---- kernel-level entry points for parity tests (mmc_dbg_*) -------------------------------- */

/* ComputePairInteractionEnergy_singlemol (energy_utils.f90:374) at the current geometry */
int mmc_dbg_pair_singlemol(mmc_t* handle, int chain, int res, int mol, double* e_non_coulomb, double* e_coulomb);
/* ComputeEwaldSelfInteractionSingleMol (ewald_energy.f90:308) */
int mmc_dbg_self_singlemol(mmc_t* handle, int res, double* e_self);
/* ComputeIntraResidueRealCoulombEnergySingleMol (ewald_energy.f90:371) */
int mmc_dbg_intra_singlemol(mmc_t* handle, int chain, int res, int mol, double* e_intra);
/* LookupTabulated (tabulated_utils.f90:87) evaluated on the device: which = 0 erfc(alpha r)/r, 1 r^6, 2 r^12 */
int mmc_dbg_table_lookup(mmc_t* handle, int which, const double* r, int n, double* out);
/* device erfc used by the pair sweep, for accuracy tests */
int mmc_dbg_erfc(mmc_t* handle, const double* x, int n, double* out);
/* Allocation-free variant of mmc_dbg_erfc.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same kernel and same results as mmc_dbg_erfc, but it calls no cudaMalloc / cudaFree / allocating routine: the
 * device staging scratch is one file-scope device array of MMC_NOALLOC_MAX_N inputs + MMC_NOALLOC_MAX_N outputs (1 MiB) that
 * every GPU loading the library holds until exit.  x and out are host buffers of n doubles owned by the caller.
 * n > MMC_NOALLOC_MAX_N is rejected with an error.  All calls on one GPU share the scratch: it keeps the last call's contents
 * (nothing needs clearing: every slot that is read was written by this call first), and calls that overlap on the same GPU
 * race on it -- serialise them.  Returns after its own work has drained. */
#define MMC_NOALLOC_MAX_N 65536
int mmc_erfc_noalloc(mmc_t* handle, const double* x, int n, double* out);
/* measurement helper (bench.py strict_mode): n_moves translation moves of one chain driven from the host the way the
 * Fortran Translation routine drives them, through mmc_trial_step (fused = 1: one launch per move) or through
 * mmc_trial_energy + mmc_trial_commit (fused = 0); wall seconds and accepted moves */
int mmc_dbg_strict_loop(mmc_t* handle, int chain, int n_moves, uint64_t seed, int fused, double* seconds, int64_t* n_accepted);
/* test support: overwrite a chain's tracked energies (6 doubles) and / or tracked A(k) (2*K doubles); NULL = keep */
int mmc_dbg_set_tracked(mmc_t* handle, int chain, const double* energies, const double* sk);
/* timing of the last pair / recip sweeps in a batched run is taken by the caller with
 * CUDA events on this stream */
void* mmc_stream(mmc_t* handle);
/* launch configuration of the batched kernel chosen for this handle: CTA size and CTAs per chain (1 = one CTA owns a chain;
 * 2 / 4 / 8 = a thread-block cluster shares it, chosen when chains <= SMs so that few chains still fill the GPU;
 * environment variable MANIAC_CLUSTER=<0|2|4|8> overrides the choice at mmc_init) */
int mmc_chain_launch_config(const mmc_t* handle, int* threads, int* cluster);
/* kernel launches issued by this handle since init (bench.py's gpu_launches) */
int64_t mmc_launch_count(const mmc_t* handle);
/* Executed real-space work of the per-move sweeps since mmc_init (all chains): out[0] = minimum-image distances
 * formed, out[1] = pairs found inside the reach (Coulomb kernel and/or LJ evaluated).  The reference loop
 * (energy_utils.f90:399-437) forms every distance; the tile-culled sweep forms only these.  Used by bench.py for
 * the executed-flop figure next to the algorithmic one. */
int mmc_pair_work(mmc_t* handle, int64_t out[2]);
/* Instrumented replay for the roofline report: runs n_steps batched steps WITHOUT graph capture,
 * CUDA events around every launch on the handle's stream, and returns the average duration in ms
 * of one launch of {pair sweep, reciprocal sweep, step tail} in ms[0..2] and the number of
 * launches of each in n[0..2].  Advances the chains like mmc_run_batch_async. */
int mmc_profile_kernels(mmc_t* handle, int n_steps, uint64_t seed, double ms[3], int64_t n[3]);
/* diagnostic builds (-DMMC_PHASE_TIMING) only: clock cycles per phase of the batched kernel, summed over CTAs and trials:
 * out[0..6] = propose, stage, cull + real-space sweep, block sum, reciprocal sweep, block sum, finalize + commit;
 * out[7] = trials counted.  All zero in production builds. */
int mmc_dbg_phase_cycles(mmc_t* handle, int64_t out[8]);
/* The framework-energy cache of one chain (no counterpart in the reference, which re-sums the framework part of the old
 * geometry's pair energy in every move, ComputeOldEnergy monte_carlo_utils.f90:347-396): out[2 * slot + {0, 1}] = {LJ,
 * Coulomb before the 1/(4 pi eps0) factor} of the guest molecule in device slot `slot` with the framework, for the g_mols
 * slots of mmc_chain_capacity; NaN where nothing is cached.  Tests only. */
int mmc_dbg_fw_cache(mmc_t* handle, int chain, double* out);
/* FP64 FMA peak micro-benchmark (register-resident DFMA chains): returns TFLOP/s */
int mmc_measure_fp64_peak(int device, double* tflops);

#ifdef __cplusplus
}
#endif
#endif /* MANIAC_GPU_H */
