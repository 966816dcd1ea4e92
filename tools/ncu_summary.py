#!/usr/bin/env python
"""Condense an `ncu --set full` report into the handful of per-launch numbers DESIGN.md / bench.py quote.

    python tools/ncu_summary.py gpurun_out/<tag>_prof.ncu-rep > profiles/<tag>_ncu_summary.txt
"""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs/thread"),
    ("launch__occupancy_limit_registers", "occ limit regs (blocks)"),
    ("launch__occupancy_limit_shared_mem", "occ limit smem (blocks)"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "FP64 pipe % of peak (active)"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "XU pipe %"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "ALU pipe %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "LSU pipe %"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "active threads / warp inst"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("lts__t_bytes.sum", "L2 bytes"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit rate %"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
    # ---- the load/store data pipe (shared + global/local data wavefronts): the limiter candidate DESIGN.md section 4 names
    ("l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed", "LSU data-pipe wavefronts % of peak"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "  of which shared memory %"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "shared-memory wavefronts"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum", "  shared loads (wavefronts)"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum", "  shared stores (wavefronts)"),
    ("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "issue slots busy %"),
    ("smsp__issue_active.avg.per_cycle_active", "IPC per scheduler"),
    # ---- instruction supply (code size): L0/L1 instruction caches and the GPC-level cache
    ("sm__icc_request_hit_rate.pct", "ICC (L1 instruction) hit rate %"),
    ("sm__icc_requests.sum", "ICC requests"),
    ("gcc__cache_requests_type_instruction.sum", "GCC instruction requests (ICC misses)"),
    ("gcc__average_cache_request_hit_rate.pct", "GCC hit rate %"),
    ("idc__request_hit_rate.pct", "IDC (constant) hit rate %"),
    # ---- local memory (spills / local arrays)
    ("sass__inst_executed_local_loads", "local loads executed (warp inst)"),
    ("sass__inst_executed_local_stores", "local stores executed (warp inst)"),
    ("l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum", "local load sectors"),
    ("l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum", "local store sectors"),
    ("smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "stall long scoreboard"),
    ("smsp__average_warp_latency_issue_stalled_short_scoreboard.ratio", "stall short scoreboard"),
    ("smsp__average_warp_latency_issue_stalled_math_pipe_throttle.ratio", "stall math pipe throttle"),
    ("smsp__average_warp_latency_issue_stalled_barrier.ratio", "stall barrier"),
    ("smsp__average_warp_latency_issue_stalled_wait.ratio", "stall wait"),
    ("smsp__average_warp_latency_issue_stalled_mio_throttle.ratio", "stall mio throttle"),
    ("smsp__average_warp_latency_issue_stalled_lg_throttle.ratio", "stall lg throttle"),
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    print(f"# {rep}: {len(data)} captured launches (ncu --set full --clock-control none)")
    for r in data:
        print(f"\n== {r[col['Kernel Name']]}  (launch id {r[col['ID']]})")
        for key, label in KEYS:
            if key in col:
                print(f"  {label:34s} {r[col[key]]:>16s} {units[col[key]]}")
        if "dram__bytes_read.sum" in col:
            def mb(k):
                v, u = float(r[col[k]]), units[col[k]]
                return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
            print(f"  {'DRAM traffic (read+write) bytes':34s} {mb('dram__bytes_read.sum') + mb('dram__bytes_write.sum'):16.0f}")
        # complete warp-state mix from the PC sampler (every class, so that the list sums to 100 %)
        pcs = {h[len("smsp__pcsamp_warps_issue_stalled_"):]: float(r[i] or 0) for h, i in col.items()
               if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued")}
        tot = sum(pcs.values())
        if tot > 0:
            print("  warp-state samples (all classes, %% of %d):" % int(tot))
            print("    " + ", ".join(f"{k} {100 * v / tot:.1f}" for k, v in sorted(pcs.items(), key=lambda x: -x[1]) if v > 0))


if __name__ == "__main__":
    main()
