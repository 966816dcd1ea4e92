// kernels_chain.cuh -- batched mode: the inner `do current_step` loop of MonteCarloLoop
// (monte_carlo.f90:42-76) for MANY independent chains (pressure points x replicas) in one launch.
//
// Chains never exchange data, so a chain needs no grid-wide synchronisation: ONE CTA owns a chain for the
// whole launch and loops over the MC steps; block barriers are the only synchronisation.  Per step:
//   thread 0      : selection + proposal (propose_move, draw order of SURVEY 3.8)
//   all threads   : stage the trial in shared memory (positions, LJ rows, 1-D phase tables)
//   all threads   : real-space sweep over the framework tiles in reach + this chain's guests (FP64 pipe)
//   all threads   : reciprocal sweep A'(k) = A(k) + dS(k), sum f W |A|^2 (HBM / shared memory)
//   thread 0      : energy assembly, Metropolis test, commit (coordinates, A buffer flip), record
// Several CTAs are resident per SM, so one chain's serial sections overlap other chains' sweeps.
// grid = chains handled by this launch (chain = chain0 + blockIdx.x).
#pragma once
#include <cooperative_groups.h>
#include "kernels_move.cuh"
#include "kernels_sweep.cuh"

// phases: bit 0 = real-space sweep, bit 1 = reciprocal sweep (both for production; single-phase variants
// exist only so bench.py can time each sweep alone on the same trials -- they do not commit)
template <int NT, int MINB, bool TRI>
__global__ void __launch_bounds__(NT, MINB) k_chain_run(DevConst dc, ChainState cs, const BatchCtl* ctl, int chain0, int n_steps,
                                                        int phases) {
    extern __shared__ __align__(16) unsigned char smem[];
    const SweepCtx s = sweep_carve(dc, smem, NT);
    const int c = chain0 + blockIdx.x;
    sweep_load_tables(dc, s);
    pair_count_init(s);
    for (int st = 0; st < n_steps; st++) {
        if (threadIdx.x < 32) {
            // the step's uniforms: Philox4x32-10 is ~120 dependent integer instructions per draw; 16 lanes compute draws
            // 0..15 side by side instead of thread 0 computing 6-10 of them one after the other (tape replay reads the tape)
            const bool philox = ctl->rng.tape == nullptr;
            if (philox && threadIdx.x < MMC_PRE_DRAWS)
                s.mol->u[threadIdx.x] = philox_uniform(ctl->rng.seed, cs.chain_id[c], ctl->rng.step + (unsigned long long)st, threadIdx.x);
            __syncwarp();
            if (threadIdx.x == 0) {   // the RNG source is re-read here: it must not occupy registers during the sweeps
                RngSrc rs = ctl->rng;
                rs.step += (unsigned long long)st;
                propose_move(dc, cs, rs, c, *s.trial, philox ? s.mol->u : nullptr);
                if (s.trial->error) atomicOr(cs.error_flag, MMC_ERRF_CAPACITY);
            }
        }
        __syncthreads();
        MMC_TSTAMP(0);
        sweep_trial<NT, TRI>(dc, cs, c, s, *s.trial, 0, 1, phases, true);
        if (threadIdx.x == 0) {
            Trial& t = *s.trial;   // shared memory: the single-thread sections no longer round-trip ~2 kB through global memory
            const int kind = t.kind;
            if (kind > 0) {
                fold_fw_cache(t, *s.mol, s.mol->sums, s.mol->fws);
                finalize_trial(dc, cs, c, t, s.mol->sums, s.mol->intra_ok ? s.mol->intra : nullptr);
                RngSrc rs = ctl->rng;
                rs.step += (unsigned long long)st;
                Rng rng{&rs, cs, c, t.draw, rs.tape ? nullptr : s.mol->u};  // the accept draw follows the proposal's draws
                const double u = rng.next();
                const int acc = (phases == 3 && u <= t.prob) ? 1 : 0;
                // AcceptDeletionMove draws three more uniforms to place the molecule in the reservoir box (delete_molecule.f90:150-151)
                if (acc && kind == MMC_DELETION && dc.has_reservoir) { rng.next(); rng.next(); rng.next(); }
                commit_trial(dc, cs, c, t, acc);
            }
            if (ctl->recs) {
                mmc_move_record& rec = ctl->recs[(size_t)c * ctl->rec_stride + ctl->rec_base + st];
                rec.kind = kind > 0 ? kind : 0; rec.res = kind > 0 ? (kind == MMC_SWAP ? t.res + 256 * t.res2 : t.res) : 0; rec.mol = kind > 0 ? t.mol : 0;
                rec.accepted = kind > 0 ? t.accepted : 0; rec.prob = kind > 0 ? t.prob : 0.0;
                for (int i = 0; i < MMC_NE; i++) { rec.old_e[i] = kind > 0 ? t.old_e[i] : 0.0; rec.new_e[i] = kind > 0 ? t.new_e[i] : 0.0; }
            }
        }
        __syncthreads();
#ifdef MMC_PHASE_TIMING
        if (threadIdx.x == 0) s.mol->phase[7] += 1ull;
#endif
        MMC_TSTAMP(6);
    }
    pair_count_flush(cs, s);
}

// ---- few chains: a thread-block CLUSTER per chain ---------------------------------------------------------------------
// With fewer chains than the GPU has room for (the north star's own 8-GPU isotherm: 512 chains / 8 = 64 chains on 148 SMs)
// one CTA per chain leaves SMs idle and every chain latency-bound.  Here P = cluster size CTAs share a chain:
//   every CTA   : the SAME proposal from the same Philox counters into its own shared-memory Trial (only rank 0 performs the
//                 proposal's side effects: counters, slot geometry of a creation), then 1/P of the tiles and of the (kx, ky)
//                 rows (sweep_trial with part = cluster rank, exactly what k_sweep does across CTAs in strict mode)
//   cluster.sync: the six block sums of the P CTAs are complete
//   rank 0      : adds them in rank order through distributed shared memory, assembles the energies, draws the accept uniform,
//                 commits (coordinates, slot compaction, A buffer flip), writes the record
//   cluster.sync: (release / acquire at cluster scope) the committed state is visible to the peers' next proposal
// No host round trip, no extra launch; block barriers inside the CTAs as before.  Philox streams only (a tape is consumed
// through a global cursor: tape replay runs the one-CTA kernel).  The records equal the one-CTA kernel's up to the
// association of the P partial sums (1e-16 relative); the accept sequence is identical.
// grid = P x chains, cluster dimension P (launch attribute); chain = chain0 + blockIdx.x / P.  Two CTAs of 512 threads per SM
// (64 registers): without the minimum in __launch_bounds__ ptxas took 128 registers = ONE CTA per SM and every grid above 148
// CTAs ran in two waves (r2m: 64 chains x P = 4 at 45 us per step against 22.5 us for 32 chains).
template <int NT, bool TRI>
__global__ void __launch_bounds__(NT, 2048 / NT / 2) k_chain_run_cl(DevConst dc, ChainState cs, const BatchCtl* ctl, int chain0, int n_steps, int phases) {
    namespace cg = cooperative_groups;
    extern __shared__ __align__(16) unsigned char smem[];
    cg::cluster_group cluster = cg::this_cluster();
    const int P = (int)cluster.num_blocks(), rank = (int)cluster.block_rank();
    const bool lead = rank == 0;
    const SweepCtx s = sweep_carve(dc, smem, NT);
    const int c = chain0 + blockIdx.x / P;
    sweep_load_tables(dc, s);
    pair_count_init(s);
    for (int st = 0; st < n_steps; st++) {
        if (threadIdx.x < 32) {
            if (threadIdx.x < MMC_PRE_DRAWS)
                s.mol->u[threadIdx.x] = philox_uniform(ctl->rng.seed, cs.chain_id[c], ctl->rng.step + (unsigned long long)st, threadIdx.x);
            __syncwarp();
            if (threadIdx.x == 0) {
                RngSrc rs = ctl->rng;
                rs.tape = nullptr;
                rs.step += (unsigned long long)st;
                propose_move(dc, cs, rs, c, *s.trial, s.mol->u, lead);
                if (lead && s.trial->error) atomicOr(cs.error_flag, MMC_ERRF_CAPACITY);
            }
        }
        __syncthreads();
        sweep_trial<NT, TRI>(dc, cs, c, s, *s.trial, rank, P, phases, true);
        cluster.sync();                                   // every CTA's s.mol->sums is final
        if (lead && threadIdx.x == 0) {
            Trial& t = *s.trial;
            const int kind = t.kind;
            if (kind > 0) {
                double v[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0}, fw[4] = {0.0, 0.0, 0.0, 0.0};
                for (int r = 0; r < P; r++) {             // fixed order: deterministic
                    const double* ps = cluster.map_shared_rank(s.mol->sums, r);
                    for (int i = 0; i < 6; i++) v[i] += ps[i];
                    const double* pf = cluster.map_shared_rank(s.mol->fws, r);
                    for (int i = 0; i < 4; i++) fw[i] += pf[i];
                }
                fold_fw_cache(t, *s.mol, v, fw);          // the framework-energy cache (every rank took the same decision)
                finalize_trial(dc, cs, c, t, v, s.mol->intra_ok ? s.mol->intra : nullptr);
                RngSrc rs = ctl->rng;
                rs.tape = nullptr;
                rs.step += (unsigned long long)st;
                Rng rng{&rs, cs, c, t.draw, s.mol->u};    // the accept draw follows the proposal's draws
                const double u = rng.next();
                const int acc = (phases == 3 && u <= t.prob) ? 1 : 0;
                if (acc && kind == MMC_DELETION && dc.has_reservoir) { rng.next(); rng.next(); rng.next(); }
                commit_trial(dc, cs, c, t, acc);
            }
            if (ctl->recs) {
                mmc_move_record& rec = ctl->recs[(size_t)c * ctl->rec_stride + ctl->rec_base + st];
                rec.kind = kind > 0 ? kind : 0; rec.res = kind > 0 ? (kind == MMC_SWAP ? t.res + 256 * t.res2 : t.res) : 0; rec.mol = kind > 0 ? t.mol : 0;
                rec.accepted = kind > 0 ? t.accepted : 0; rec.prob = kind > 0 ? t.prob : 0.0;
                for (int i = 0; i < MMC_NE; i++) { rec.old_e[i] = kind > 0 ? t.old_e[i] : 0.0; rec.new_e[i] = kind > 0 ? t.new_e[i] : 0.0; }
            }
            __threadfence();
        }
        cluster.sync();                                   // the committed state is visible to every CTA of the cluster
    }
    pair_count_flush(cs, s);
}
