"""Per-source-line instruction and stall-sample totals from an ncu report (cuda,sass source view).
usage: python tools/hot_lines.py report.ncu-rep [warps_per_launch] [samples]"""
import csv, subprocess, sys
rep = sys.argv[1]
div = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
fname = kernel = None
hdr = None
agg = {}
tot = 0
for r in rows:
    if not r: continue
    if r[0] == "File Path": fname = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": kernel = r[1]; continue
    if r[0] == "Line No": hdr = r; iN = hdr.index("Instructions Executed"); iS = hdr.index("# Samples"); continue
    if hdr is None: continue
    if r[0] != "":  # a source line row: carries the sum over its sass
        try: n = int(r[iN]); s = int(r[iS])
        except ValueError: continue
        if n == 0 and s == 0: continue
        key = (fname, int(r[0]))
        a = agg.setdefault(key, [0, 0, r[1].strip()[:110]])
        a[0] += n; a[1] += s
        tot += n
print("kernel", kernel, "total inst", tot, "per warp %.1f" % (tot / div))
by = 1 if len(sys.argv) > 3 and sys.argv[3] == "samples" else 0
print("total samples", sum(v[1] for v in agg.values()))
for (f, l), (n, s, src) in sorted(agg.items(), key=lambda kv: -kv[1][by])[:70]:
    print("%-16s:%4d inst/warp=%7.1f (%4.1f%%) samples=%5d  %s" % (f, l, n / div, 100.0 * n / tot, s, src))
