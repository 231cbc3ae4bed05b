"""Per-source-line instruction and stall-sample shares from an .ncu-rep (needs -lineinfo).
usage: python tools/ncu_lines.py x.ncu-rep [topN]"""
import collections, csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
agg = collections.defaultdict(lambda: [0.0, 0.0, ""])
cur, hdr = None, None
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        hdr = r; si = hdr.index("# Samples"); ii = hdr.index("Instructions Executed"); continue
    if hdr is None or len(r) <= max(si, ii) or not r[0].strip().isdigit():
        continue
    try:
        s, ie = float(r[si] or 0), float(r[ii] or 0)
    except ValueError:
        continue
    k = (cur, int(r[0]))
    agg[k][0] += s; agg[k][1] += ie; agg[k][2] = r[1].strip()[:95]
tot = sum(v[0] for v in agg.values()) or 1
toti = sum(v[1] for v in agg.values()) or 1
print(f"total samples {tot:.0f}, warp instructions {toti:.0f}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{100*v[0]/tot:5.1f}% smp {100*v[1]/toti:5.1f}% inst  {k[0]}:{k[1]:<4d} {v[2]}")
