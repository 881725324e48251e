"""Warp-stall samples and executed instructions per CUDA source line of an .ncu-rep (needs -lineinfo and
--import-source on).  Usage: python scripts/ncu_lines.py report.ncu-rep [top_n]"""
import csv, subprocess, sys
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = None; recs = []; fname = ""
for r in rows:
    if len(r) == 2 and r[0] == "File Path": fname = r[1].split("/")[-1]
    if r and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr) and r[2] == "-" and r[0].isdigit():  # a CUDA source line row (no SASS address)
        d = dict(zip(hdr, r)); recs.append((fname, int(r[0]), r[1], int(d["# Samples"] or 0), int(d["Instructions Executed"] or 0)))
tot = sum(x[3] for x in recs); toti = sum(x[4] for x in recs)
print(f"total samples {tot}, warp instructions {toti}")
for f, ln, src, smp, ins in sorted(recs, key=lambda x: -x[3])[:topn]:
    print(f"{smp:6d} {100.0 * smp / max(tot, 1):5.1f}%  inst {100.0 * ins / max(toti, 1):5.1f}%  {f}:{ln}: {src.strip()[:110]}")
