#!/usr/bin/env python
"""bench.py -- MC trial moves/s of the per-move energy path on B200 (BASELINE.json metric), with the
roofline of the dominant kernel and the CPU restatement timed beside it.

    python bench.py --gpus N --steps K --warmup W            # own arm (CUDA library through the C ABI)
    python bench.py --impl reference --gpus N --steps K ...  # reference arm: the CPU oracle port on the host cores

Workload (config.workload = "c4_large_k"): BASELINE.json configs[3] -- synthetic 50 000-atom framework (one
inactive residue) + 200 rigid TIP4P guests per chain, L = 96 A, r_c = 12 A, Ewald tol 1e-6 -> K = 37 782
k-vectors; `--chains` independent chains per GPU (isotherm: 64 log-spaced fugacities x replicas, configs[4]),
`--moves` MC steps per chain per bench step, translation/rotation/insertion/deletion 0.4/0.4/0.1/0.1.
One bench "step" = one batch: chains x moves trial moves (chain-per-CTA kernel k_chain_run, 128 MC steps per launch,
launches replayed from a CUDA graph).

Multi-GPU: one process per GPU (torchrun); chains (pressure points x replicas) are dealt to ranks, no data-path
collective; weak scaling (chains per GPU fixed).  The only collectives are the barrier and the MAX of
the device-timed durations.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

W_PAIR_FLOP = 90.0   # FP64 flop per pair-eval, SURVEY.md 8(d) / DESIGN.md
B_K_BYTES = 40.0     # HBM bytes per k-eval: A(k) read 16 + A'(k) write 16 + f*W 8
FUG_LO_ATM, FUG_HI_ATM, N_POINTS = 1.0e-2, 1.0e4, 64
# dram__bytes_read.sum + dram__bytes_write.sum of one k_chain_run launch from the committed `ncu --set full` capture,
# divided by that launch's k-evals (592 chains x 32 steps x 37 782 k); bench.py scales it to its own launch.
NCU_DRAM_BYTES_PER_K_EVAL = 23464590000.0 / (592 * 32 * 37782)
NCU_TRAFFIC_SOURCE = "profiles/r2zp_ncu_summary.txt (23.46 GB per launch of 592 chains x 32 steps), scaled by k-evals"


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="own", choices=["own", "reference"])
    p.add_argument("--chains", type=int, default=592, help="independent chains per GPU (4 resident CTAs x 148 SMs)")
    p.add_argument("--moves", type=int, default=256, help="MC steps per chain per bench step")
    p.add_argument("--framework", type=int, default=50000)
    p.add_argument("--guests", type=int, default=200)
    p.add_argument("--cpu-moves", type=int, default=100, help="moves per host thread per reference-arm step; the cpu_baseline sample is 4x that")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-extras", action="store_true", help="skip the strict_mode and c5_isotherm objects")
    return p.parse_args()


def make_system(args):
    from maniac_mc_b200 import builders
    s = builders.synthetic_framework(n_framework=args.framework, n_guests=args.guests, box=96.0, cutoff=12.0,
                                     tol=1e-6, seed=4, capacity=512, p_translation=0.4, p_rotation=0.4)
    s.init_sk, s.deletion_mode = 1, 1   # physically consistent A(k) (quirk switches Q1/Q2 off); work per move is unchanged
    return s


def chain_fugacity(s, global_chain):
    from maniac_mc_b200 import builders
    point = global_chain % N_POINTS
    f_atm = FUG_LO_ATM * (FUG_HI_ATM / FUG_LO_ATM) ** (point / (N_POINTS - 1))
    f = s.fugacity.copy()
    f[0] = builders.fugacity_from_atm(f_atm, s.temperature)
    return f


def workload_config(args, s, world):
    """ONE description printed by BOTH arms (the driver compares them): every string is true of the device arm and of the CPU
    arm.  The two differences between the arms are spelled out here instead of hidden: the CPU chains skip the O(N K)
    single-point start-up (A(k) starts at zero, the reference's own default, quirk Q1 -- per-move work identical), and the
    CPU arm runs as many chains as the host has threads, at the isotherm's first fugacity points."""
    return {"workload": "c4_large_k", "framework_atoms": int(args.framework), "guests_per_chain_initial": int(args.guests),
            "guest_atoms_per_molecule": 4, "box_A": 96.0, "cutoff_A": 12.0, "ewald_tol": 1e-06,
            "k_vectors": int(s.num_kvectors), "chains_per_gpu": int(args.chains), "moves_per_chain_per_step": int(args.moves),
            "move_mix": "translation 0.4 / rotation 0.4 / insertion 0.1 / deletion 0.1",
            "isotherm": f"{N_POINTS} fugacities {FUG_LO_ATM:g}..{FUG_HI_ATM:g} atm x replicas, chains dealt to ranks (CPU arm: one chain "
                        "per host thread at fugacity points 0..threads-1, a bounded sample of the same workload)",
            "parallelism": f"chains x{world} (no data-path collective)",
            "quirks": "deletion=physical; init_sk=full on the device arm, init_sk=zero (reference default: no O(N K) start-up) on the CPU arm -- "
                      "the per-move work is the same",
            "rng": "Philox4x32-10 keyed by (seed; step, global chain id, draw), same generator in both arms",
            "real_space": "device arm: the framework part of a guest molecule's pair energy is kept per molecule (the framework never "
                          "moves) and reused as the old geometry's, everything else is summed every move; CPU arm: every pair re-summed "
                          "every move as in the reference -- same energies to 1e-10 (tests/test_gpu_parity.py::test_framework_energy_cache); "
                          "MANIAC_FW_CACHE=0 switches the device arm to the reference's re-summation",
            "l2": "256 MiB write between steps (inside the timed region, device arm)"}


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.proc = None; self.lines = []
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True); self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def pair_evals_of(s, args, counts, d_counters):
    """pair-evals done by the moves counted in d_counters (per chain dicts of counter deltas)."""
    nat, n_fw = 4, args.framework
    total = 0.0
    for c, d in enumerate(d_counters):
        n_atoms = n_fw + nat * float(counts[c])
        total += (d["trial_translations"] + d["trial_rotations"]) * 2.0 * nat * (n_atoms - nat)
        total += d["trial_creations"] * nat * n_atoms + d["trial_deletions"] * nat * (n_atoms - nat)
    return total


def counters_delta(a, b):
    return [{k: b[i][k] - a[i][k] for k in b[i]} for i in range(len(b))]


# ------------------------------------------------------------------------------------------------
# CPU leg: the oracle port (oracle/maniac_oracle.c, plain C, glibc) -- the only place bench.py executes oracle/.
# The reference itself is pure Fortran and neither this image nor the GPU box has a Fortran compiler, so the
# baseline kind is "port".  One independent chain per host thread (the reference is single threaded; replicas
# are the only way it can use more cores).
class CpuSample:
    def __init__(self, s, n_threads):
        from oracle.oracle import Oracle
        cpu = s.copy()
        cpu.init_sk = 0   # reference default (Q1): skips the ~30 s single-point start-up per chain; per-move work identical
        self.oracles = [Oracle(cpu) for _ in range(n_threads)]
        for i, o in enumerate(self.oracles):   # the isotherm's fugacity points of chains 0 .. n_threads-1, as the GPU arm's chains
            o.set_params(chain_fugacity(s, i), s.temperature)
            o.prepare_fourier()                # every molecule's phase tables (ComputeAllFourierTerms) without the O(N K) sums
        self.step0 = 0

    def run(self, moves):
        """every thread advances its chain by `moves` MC steps; returns wall seconds"""
        def work(i, o):
            o.run(moves, seed=99, chain=i, step0=self.step0, records=False)
        ths = [threading.Thread(target=work, args=(i, o)) for i, o in enumerate(self.oracles)]
        t0 = time.perf_counter()
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        dt = time.perf_counter() - t0
        self.step0 += moves
        return dt


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def run_reference(args, rank, world):
    if rank != 0:
        return
    s = make_system(args)
    nt = host_threads()
    cs = CpuSample(s, nt)
    for _ in range(args.warmup):
        cs.run(max(1, args.cpu_moves // 4))
    t = 0.0
    for _ in range(args.steps):
        t += cs.run(args.cpu_moves)
    moves = nt * args.cpu_moves * args.steps
    v = moves / t
    sample = (f"{nt} host threads x {args.cpu_moves} MC moves per step of the same c4_large_k system, one independent chain "
              f"per thread at the isotherm's first {nt} fugacity points (oracle port, gcc -O2, reference-default init_sk=zero)")
    cfg = workload_config(args, s, world)
    print(json.dumps({"impl": "reference", "metric": "mc_trial_moves_per_s", "value": v, "unit": "moves/s", "n_gpus": args.gpus,
                      "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True,
                      "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
                      "cpu_baseline": {"value": v, "unit": "moves/s", "cores": nt, "kind": "port", "sample": sample},
                      "e2e": {"value": v, "unit": "moves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                      "reference_run": {"cpu_chains": nt, "moves_per_chain_per_step": int(args.cpu_moves), "init_sk": "zero",
                                        "fugacity_points": f"0..{min(nt, N_POINTS) - 1} of {N_POINTS}"},
                      "gpu_launches": 0}))


# ------------------------------------------------------------------------------------------------
def c5_isotherm(rank, world, local_rank, steps=1000, reps=5):
    """BASELINE configs[4], the FIXED-size sweep of the north star: 64 fugacity points x 8 replicas = 512 chains of the
    reference's ZIF-8 + H2O example IN TOTAL, dealt round-robin to the ranks (strong scaling: 512 / world chains per GPU).
    Device-timed on the library's stream, MAX over ranks.  Returns the object for the JSON line (rank 0) or None."""
    import torch
    import torch.distributed as dist
    from maniac_mc_b200 import builders, isotherm as iso
    from maniac_mc_b200.lib import ManiacGPU
    from maniac_mc_b200.system import System
    s = System.load(os.path.join(ROOT, "tests", "golden", "zif8_h2o.npz"))
    s.init_sk, s.deletion_mode = 1, 1
    n_points, replicas = 64, 8
    ids = iso.deal_chains(n_points, replicas, world, rank)
    f_atm = iso.fugacity_points(n_points, FUG_LO_ATM, FUG_HI_ATM)
    res = int(np.flatnonzero(s.is_active)[0])
    gpu = ManiacGPU(s, n_chains=len(ids), device=local_rank)
    for c, cid in enumerate(ids):
        f = s.fugacity.copy(); f[res] = builders.fugacity_from_atm(f_atm[cid.point], s.temperature)
        gpu.set_chain_params(c, f, s.temperature); gpu.set_chain_id(c, cid.seed)
    gpu.system_energy_all()
    stream = torch.cuda.ExternalStream(gpu.stream(), device=local_rank)
    for _ in range(2):
        gpu.run_batch_async(steps, seed=5)
    gpu.synchronize()
    if world > 1:
        dist.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(reps):
        gpu.run_batch_async(steps, seed=5)
    ev1.record(stream)
    gpu.synchronize()
    ms = ev0.elapsed_time(ev1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=f"cuda:{local_rank}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t[0])
    threads = list(gpu.chain_launch_config())   # (CTA threads, CTAs per chain: > 1 = the cluster-per-chain kernel)
    gpu.close()
    if rank != 0:
        return None
    total = n_points * replicas
    return {"workload": "c5_isotherm: 64 fugacity points x 8 replicas = 512 chains of ZIF-8 + H2O (2208-atom framework, K = 297) in TOTAL, "
                        "dealt round-robin to the ranks", "scaling": "strong", "chains_total": total, "chains_per_gpu": len(ids),
            "steps_per_chain": steps * reps, "value": total * steps * reps / (ms * 1e-3), "unit": "moves/s", "ms": ms,
            "launch_config_threads_ctas_per_chain": threads}


def strict_mode(gpu_c4, s_c4, local_rank, cpu_moves_per_s_per_core):
    """The drop-in (strict) mode the unmodified Fortran move routines call: one host-driven translation move per call,
    mmc_trial_step (one launch per move) and the two-call form mmc_trial_energy + mmc_trial_commit, timed by wall clock
    around a C loop inside the library (mmc_dbg_strict_loop) on the bench system and on the reference's ZIF-8 example."""
    from maniac_mc_b200.lib import ManiacGPU
    from maniac_mc_b200.system import System
    from oracle.oracle import Oracle
    out = {}
    z = System.load(os.path.join(ROOT, "tests", "golden", "zif8_h2o.npz")); z.init_sk, z.deletion_mode = 1, 1
    gz = ManiacGPU(z, n_chains=1, device=local_rank); gz.system_energy_all()
    o = Oracle(z); o.system_energy()
    t0 = time.perf_counter(); o.run(4000, seed=3, chain=0, step0=0, records=False); cpu_z = 4000 / (time.perf_counter() - t0)
    for name, g, n, cpu in (("zif8_h2o", gz, 4000, cpu_z), ("c4_large_k", gpu_c4, 2000, cpu_moves_per_s_per_core)):
        g.strict_loop(200, seed=1, fused=True)                     # warm-up
        tf, af, lf = g.strict_loop(n, seed=2, fused=True)
        t2, a2, l2 = g.strict_loop(n, seed=2, fused=False)
        out[name] = {"moves_per_s": n / tf, "us_per_move": 1e6 * tf / n, "launches_per_move": lf / n, "accepted": af,
                     "entry": "mmc_trial_step (energies + accept test + commit in one launch, result through host-mapped memory)",
                     "two_call_form": {"moves_per_s": n / t2, "us_per_move": 1e6 * t2 / n, "launches_per_move": l2 / n,
                                       "entry": "mmc_trial_energy + mmc_trial_commit"},
                     "cpu_port_single_chain_moves_per_s": cpu, "moves": n, "move_type": "translation (host proposal, host uniform)"}
    gz.close()
    return out


def run_own(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from maniac_mc_b200 import lib
    from maniac_mc_b200.lib import ManiacGPU
    from maniac_mc_b200.system import N_COUNTERS
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    s = make_system(args)
    C = args.chains
    gpu = ManiacGPU(s, n_chains=C, device=local_rank)
    for c in range(C):
        gpu.set_chain_params(c, chain_fugacity(s, rank * C + c), s.temperature)
        gpu.set_chain_id(c, rank * C + c)
    t_start = time.perf_counter()
    gpu.system_energy_all()     # ComputeSystemEnergy of every chain: the framework's S(k) once per handle, guests per chain
    startup_s = time.perf_counter() - t_start
    stream = torch.cuda.ExternalStream(gpu.stream(), device=local_rank)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local_rank}")
    g_mols, g_atoms = gpu.chain_capacity()
    pin = lambda *shape, dt=torch.float64: torch.zeros(*shape, dtype=dt).pin_memory()
    h_cnt, h_com, h_off = pin(C, s.nres, dt=torch.int32), pin(C, g_mols, 3), pin(C, g_atoms, 3)
    h_e, h_ctr = pin(C, 6), pin(C, N_COUNTERS, dt=torch.int64)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    warnings = []

    def download():
        try:
            gpu.download_chains(h_cnt.data_ptr(), h_com.data_ptr(), h_off.data_ptr(), h_e.data_ptr(), h_ctr.data_ptr())
        except lib.ManiacGpuError as e:   # the sticky device error flag (slot capacity reached): the arrays are still filled
            if str(e) not in warnings:
                warnings.append(str(e))

    def counters():
        download()
        names = ["translations", "rotations", "creations", "deletions", "trial_translations", "trial_rotations",
                 "trial_creations", "trial_deletions", "swaps", "trial_swaps"]
        a = h_ctr.numpy().copy()
        return [dict(zip(names, (int(v) for v in row))) for row in a]

    def device_step(seed):
        with torch.cuda.stream(stream):
            flush.fill_(1)
        gpu.run_batch_async(args.moves, seed=seed)

    def e2e_step(seed):
        gpu.upload_chains(h_cnt.data_ptr(), h_com.data_ptr(), h_off.data_ptr())   # H2D of the step's inputs (pinned)
        gpu.run_batch_async(args.moves, seed=seed)
        download()                                                                # D2H of the step's results (pinned)

    # ---- device-resident throughput --------------------------------------------------------------
    for w in range(args.warmup):
        device_step(7)
    gpu.synchronize()
    c0 = counters(); l0 = gpu.launch_count()
    clocks = ClockSampler(local_rank) if rank == 0 else None
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for k in range(args.steps):
        device_step(7)
    ev1.record(stream)
    gpu.synchronize()
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = gpu.launch_count() - l0
    c1 = counters()
    clk = clocks.stop() if clocks else None
    # ---- end to end through the C ABI with host buffers ------------------------------------------
    download()
    for w in range(2):
        e2e_step(11)
    barrier()
    ev2, ev3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev2.record(stream)
    for k in range(args.steps):
        e2e_step(11)
    ev3.record(stream)
    gpu.synchronize()
    barrier()
    ms_e2e = ev2.elapsed_time(ev3)
    if world > 1:
        t = torch.tensor([ms, ms_e2e], dtype=torch.float64, device=f"cuda:{local_rank}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e = float(t[0]), float(t[1])
    moves_total = world * C * args.moves * args.steps
    value = moves_total / (ms * 1e-3)
    e2e_value = moves_total / (ms_e2e * 1e-3)
    h2d = h_cnt.numel() * 4 + (h_com.numel() + h_off.numel()) * 8
    d2h = h2d + h_e.numel() * 8 + h_ctr.numel() * 8
    if rank != 0:
        gpu.close()
        if not args.no_extras:
            c5_isotherm(rank, world, local_rank)
        return
    # ---- roofline of the dominant kernel (rank 0): instrumented replay right after the timed region.
    # mmc_profile_kernels runs the SAME 32 proposals of every chain three times with CUDA events on the library's
    # stream: real-space phase only, reciprocal phase only (no commit), then the full k_chain_run launch.
    PROF_STEPS = 32
    cp0 = counters()
    counts0 = h_cnt.numpy()[:, 0].copy()
    pw0 = gpu.pair_work()
    prof = gpu.profile_kernels(PROF_STEPS, seed=13)
    pw1 = gpu.pair_work()
    cp1 = counters()
    counts = h_cnt.numpy()[:, 0].copy()
    dprof = counters_delta(cp0, cp1)
    evals = pair_evals_of(s, args, 0.5 * (counts0 + counts), dprof)
    nat = 4
    k_evals = sum(sum(d[k] for k in d if k.startswith("trial_")) for d in dprof) * float(gpu.nk)
    k_flop = sum((d["trial_translations"] + d["trial_rotations"]) * (30.0 * nat + 5.0) +
                 (d["trial_creations"] + d["trial_deletions"]) * (15.0 * nat + 5.0) for d in dprof) * float(gpu.nk)
    fp64_peak = lib.measure_fp64_peak(local_rank)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    hbm_src = "MEASURED_PEAKS.json hbm_gbs (measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    full_s, pair_s, recip_s = prof["tail_ms"] * 1e-3, prof["pair_ms"] * 1e-3, prof["recip_ms"] * 1e-3
    # profile_kernels launches the real-space phase twice on the same trials (alone, then inside the full launch)
    dist_exec, in_reach = ((b - a) / 2.0 for a, b in zip(pw0, pw1))
    # Dominant phase of the fused kernel = the reciprocal sweep (HBM): every trial streams its chain's A(k) in and
    # A'(k) out.  achieved = algorithmic bytes of the launch / duration of the WHOLE fused launch.
    recip_gbs = k_evals * B_K_BYTES / full_s / 1e9
    roofline = {"kernel": "k_chain_run (propose + real-space sweep + reciprocal sweep + accept/commit fused, %d MC steps per launch, "
                          "one CTA per chain)" % PROF_STEPS,
                "bound": "hbm", "achieved": recip_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": recip_gbs / hbm_peak,
                "traffic": NCU_DRAM_BYTES_PER_K_EVAL * k_evals if NCU_DRAM_BYTES_PER_K_EVAL else None,
                "traffic_source": NCU_TRAFFIC_SOURCE, "peak_source": hbm_src,
                "algorithmic": f"{B_K_BYTES:g} B per k-eval x k-evals per launch (A read 16 + A' write 16 + f*W 8; f*W is shared by all "
                               "chains and stays in L2, so DRAM traffic is ~32 B per k-eval)",
                "algorithmic_bytes_per_launch": k_evals * B_K_BYTES,
                "avg_launch_ms": prof["tail_ms"], "share_of_step": 1.0, "k_evals_per_s": k_evals / full_s,
                "phase_only_ms": {"real_space": prof["pair_ms"], "reciprocal": prof["recip_ms"], "full": prof["tail_ms"]},
                "reciprocal_phase_alone_gbs": k_evals * B_K_BYTES / recip_s / 1e9,
                "reciprocal_phase_alone_frac": k_evals * B_K_BYTES / recip_s / 1e9 / hbm_peak}
    # Real-space phase (FP64 pipe).  Two figures: the reference loop's pair-evals (every minimum-image pair, 90 flop
    # each) and the work the tile-culled sweep actually executes (distance formed: 12 flop; pair inside
    # reach: + 10-term Coulomb table 24 flop + LJ 10 flop).  The first can exceed the FP64 peak -- that is the culling.
    exec_flop = dist_exec * 12.0 + in_reach * 34.0
    roofline_pair = {"kernel": "k_chain_run, real-space share", "bound": "fp64", "unit": "TFLOP/s", "peak": fp64_peak,
                     "peak_source": "DFMA micro-benchmark run in this process (mmc_measure_fp64_peak); MEASURED_PEAKS.json has no FP64 entry",
                     "achieved": exec_flop / pair_s / 1e12, "frac": exec_flop / pair_s / 1e12 / fp64_peak,
                     "algorithmic": "executed work: 12 flop per distance formed + 34 flop per pair inside reach (counted on the device, "
                                    "mmc_pair_work); phase timed alone on the same trials",
                     "reference_loop_pair_evals_per_s": evals / pair_s,
                     "reference_loop_tflops_equivalent": evals * W_PAIR_FLOP / pair_s / 1e12,
                     "distances_formed_per_reference_pair": dist_exec / max(evals, 1.0),
                     "pairs_in_reach_per_reference_pair": in_reach / max(evals, 1.0),
                     "phase_alone_ms": prof["pair_ms"]}
    dall = counters_delta(c0, c1)
    evals_timed = pair_evals_of(s, args, counts, dall)
    out = {"metric": "mc_trial_moves_per_s", "value": value, "unit": "moves/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(args, s, world),
           "pair_evals_per_s": world * evals_timed / (ms * 1e-3),
           "e2e": {"value": e2e_value, "unit": "moves/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                   "call": "mmc_upload_chains + mmc_run_batch_async + mmc_download_chains per step, pinned host buffers"},
           "gpu_launches": int(launches), "roofline": roofline, "roofline_pair": roofline_pair,
           "clocks": clk, "mean_guests_per_chain": float(counts.mean()), "max_guests_per_chain": int(counts.max()),
           "warnings": warnings, "startup_system_energy_all_s": startup_s}
    cpu_per_core = None
    if world == 1 and not args.no_cpu_baseline:
        nt = host_threads()
        cs = CpuSample(s, nt)
        cs.run(1)
        n_cpu = 4 * args.cpu_moves
        dt = cs.run(n_cpu)
        out["cpu_baseline"] = {"value": nt * n_cpu / dt, "unit": "moves/s", "cores": nt, "kind": "port",
                               "sample": f"{nt} host threads x {n_cpu} MC moves of the same c4_large_k system, one "
                                         f"independent chain per thread (oracle port, gcc -O2, init_sk=zero), {dt:.1f} s wall"}
        cpu_per_core = nt * n_cpu / dt / nt
        del cs
    if not args.no_extras:
        try:
            out["strict_mode"] = strict_mode(gpu, s, local_rank, cpu_per_core)
        except Exception as e:   # never lose the headline line to an extra
            out["strict_mode"] = {"error": str(e)}
    gpu.close()
    if not args.no_extras:
        try:
            out["c5_isotherm"] = c5_isotherm(rank, world, local_rank)
        except Exception as e:
            out["c5_isotherm"] = {"error": str(e)}
    print(json.dumps(out))


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    run_own(args, rank, world, local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
