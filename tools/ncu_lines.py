#!/usr/bin/env python
"""Per-source-line digest of an `ncu --set full --import-source on` report: stall samples, shared-memory
wavefronts, global tag requests and executed instructions, summed over the captured launches.

    python tools/ncu_lines.py gpurun_out/<tag>_prof.ncu-rep [top_n] > profiles/<tag>_ncu_lines.txt
"""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    cur, hdr, agg = None, None, {}
    for r in csv.reader(io.StringIO(raw)):
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]; continue
        if r[0] == "Function Name":
            continue
        if r[0] == "Line No":
            hdr = r; continue
        if r[0].isdigit() and hdr:
            d = dict(zip(hdr, r))
            try:
                a = agg.setdefault((cur, int(r[0])), [0, 0, 0, 0, {}, r[1].strip()[:84]])
                a[0] += int(r[4]); a[1] += int(d["L1 Wavefronts Shared"]); a[2] += int(d["L1 Tag Requests Global"])
                a[3] += int(d["Instructions Executed"])
                for k in ("stall_barrier", "stall_long_sb", "stall_short_sb", "stall_wait", "stall_math", "stall_mio",
                          "stall_lg", "stall_not_selected", "stall_branch_resolving", "stall_dispatch"):
                    a[4][k] = a[4].get(k, 0) + int(d.get(k, 0) or 0)
            except (ValueError, KeyError):
                pass
    tot = [sum(a[i] for a in agg.values()) or 1 for i in range(4)]
    print(f"# {rep}: samples {tot[0]}, smem wavefronts {tot[1]}, global tag requests {tot[2]}, warp instructions {tot[3]} (all launches)")
    st = {}
    for a in agg.values():
        for k, v in a[4].items():
            st[k] = st.get(k, 0) + v
    print("# stall mix: " + ", ".join(f"{k[6:]} {100 * v / tot[0]:.1f}%" for k, v in sorted(st.items(), key=lambda x: -x[1])))
    for title, i in (("stall samples", 0), ("shared-memory wavefronts", 1), ("global tag requests", 2), ("warp instructions", 3)):
        print(f"\n== top lines by {title}")
        for k, a in sorted(agg.items(), key=lambda x: -x[1][i])[:top]:
            if a[i] == 0:
                break
            main_stall = max(a[4].items(), key=lambda x: x[1])[0][6:] if i == 0 and a[4] else ""
            print(f"{100 * a[i] / tot[i]:5.1f}%  {k[0]}:{k[1]:<4d} {main_stall:14s} | {a[5]}")


if __name__ == "__main__":
    main()
