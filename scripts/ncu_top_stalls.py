"""Print the SASS lines that collect most warp-stall samples in an .ncu-rep (first kernel)."""
import csv, subprocess, sys
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = [i for i, r in enumerate(rows) if 'Source' in r and any('Sampl' in c for c in r)]
h = rows[hi[0]]; si = h.index('Source'); ci = [i for i, c in enumerate(h) if 'Samples' in c][0]
body = rows[hi[0] + 1: hi[1] if len(hi) > 1 else len(rows)]
tot = sum(int(r[ci]) for r in body if len(r) > ci and r[ci].isdigit())
print("total samples", tot, "lines", len(body))
top = sorted([(int(r[ci]), n) for n, r in enumerate(body) if len(r) > ci and r[ci].isdigit()], reverse=True)[:topn]
for cnt, n in top:
    ctx = " | ".join(body[k][si].strip() for k in range(max(0, n - 2), n + 1))
    print(f"{cnt:6d} {100.0 * cnt / tot:5.1f}%  line {n}: {ctx}")
