"""Dynamic SASS instruction mix of a kernel from an .ncu-rep (source page)."""
import csv, subprocess, sys
from collections import defaultdict
rep = sys.argv[1]; units = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]; si = hdr.index("Source"); ei = hdr.index("Instructions Executed"); wi = hdr.index("Warp Stall Sampling (All Samples)")
d = defaultdict(float); st = defaultdict(float); tot = 0
for r in rows[2:]:
    if len(r) <= ei: continue
    t = r[si].split()
    if not t: continue
    op = t[1] if t[0].startswith("@") else t[0]
    n = float(r[ei]); d[op] += n; tot += n; st[op] += float(r[wi] or 0)
print("total %.0f  per-unit %.3f" % (tot, tot / units))
for k, v in sorted(d.items(), key=lambda x: -x[1])[:30]:
    print("%-28s %14.0f  %.3f/unit  stall-samples %d" % (k, v, v / units, st[k]))
