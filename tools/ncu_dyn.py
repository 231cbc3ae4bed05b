"""Dynamic warp-instruction count per source line from an .ncu-rep (development tool; needs -lineinfo).
usage: python tools/ncu_dyn.py x.ncu-rep num_warps [topN]"""
import collections, csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
nw = float(sys.argv[2]); top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
agg = collections.defaultdict(lambda: [0.0, 0.0, ""])
cur, hdr = None, None
for r in csv.reader(io.StringIO(out)):
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; si = hdr.index("# Samples"); ii = hdr.index("Instructions Executed"); continue
    if hdr is None or len(r) <= max(si, ii) or not r[0].strip().isdigit(): continue
    try: s, ie = float(r[si] or 0), float(r[ii] or 0)
    except ValueError: continue
    k = (cur, int(r[0])); agg[k][0] += s; agg[k][1] += ie; agg[k][2] = r[1].strip()[:90]
print("total inst/warp %.0f, samples %.0f" % (sum(v[1] for v in agg.values()) / nw, sum(v[0] for v in agg.values())))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{v[1]/nw:7.1f} inst/warp {v[0]:5.0f} smp {k[0]}:{k[1]:<4d} {v[2]}")
