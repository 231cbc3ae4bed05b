"""Summarise an .ncu-rep (``ncu --set full``) into the handful of metrics DESIGN.md / bench.py cite.

usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep [--src] [> profiles/x.txt]
``--src`` adds the per-source-line stall table (top lines by stall samples; needs -lineinfo at compile time).
"""
import csv
import io
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
    "launch__registers_per_thread", "launch__block_size", "launch__grid_size", "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "launch__occupancy_limit_warps",
    "launch__waves_per_multiprocessor",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "lts__t_bytes.sum", "l1tex__t_bytes.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__inst_executed_pipe_fp64.sum", "smsp__inst_executed_pipe_xu.sum", "smsp__inst_executed_pipe_lsu.sum",
    "smsp__inst_executed_pipe_fma.sum", "smsp__inst_executed_pipe_alu.sum",
]
STALL = "smsp__average_warp_latency_issue_stalled_"
STALL2 = "smsp__average_warps_issue_stalled_"


def raw(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2:]


def main():
    path = sys.argv[1]
    hdr, units, rows = raw(path)
    for r in rows:
        print("=== kernel:", r[hdr.index("Kernel Name")][:70], "| id", r[0])
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print(f"  {w:72s} {r[i]:>18s} {units[i]}")
        stalls = []
        for i, h in enumerate(hdr):
            if (h.startswith("smsp__average_warp") and "issue_stalled" in h and h.endswith("_per_issue_active.ratio")) or \
               (h.startswith("smsp__warp_issue_stalled") and h.endswith("per_warp_active.pct")):
                try:
                    stalls.append((float(r[i].replace(",", "")), h))
                except ValueError:
                    pass
        stalls.sort(reverse=True)
        for v, h in stalls[:8]:
            print(f"  stall {h:66s} {v:18.2f}")
    if "--src" in sys.argv:
        out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda"],
                             capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        # find header row
        for k, r in enumerate(rows):
            if "Source" in r and any("Sampl" in c for c in r):
                hdr = r
                body = rows[k + 1:]
                break
        else:
            print("no source page"); return
        si = hdr.index("Source")
        samp = [i for i, c in enumerate(hdr) if c.startswith("# Samples") or c == "Warp Stall Sampling (All Samples)"]
        if not samp:
            samp = [i for i, c in enumerate(hdr) if "Sampl" in c]
        ci = samp[0]
        ii = hdr.index("Instructions Executed") if "Instructions Executed" in hdr else None
        lines = []
        for r in body:
            if len(r) <= ci:
                continue
            try:
                s = float(r[ci].replace(",", "") or 0)
            except ValueError:
                continue
            lines.append((s, r[0], r[si].strip()[:110], r[ii] if ii is not None else ""))
        tot = sum(l[0] for l in lines) or 1
        print(f"=== source lines by stall samples ({hdr[ci]}), total {tot:.0f}")
        for s, ln, src, ie in sorted(lines, reverse=True)[:45]:
            print(f"  {100 * s / tot:5.1f}%  inst={ie:>12s}  L{ln:>5s}  {src}")


if __name__ == "__main__":
    main()
