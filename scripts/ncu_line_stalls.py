"""Per-source-line stall-reason breakdown of an .ncu-rep (needs -lineinfo, --import-source on) for a range of lines of
one file.  Usage: python scripts/ncu_line_stalls.py report.ncu-rep file.cu first_line last_line"""
import csv, subprocess, sys
rep, fname, lo, hi = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = None; cur = ""
stall_cols = None
for r in rows:
    if len(r) == 2 and r[0] == "File Path": cur = r[1].split("/")[-1]
    if r and r[0] == "Line No":
        hdr = r
        stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") or h.startswith("Warp Stall") or h in ("# Samples", "Instructions Executed")]
        if "--cols" in sys.argv: print(hdr)
        continue
    if hdr and cur == fname and len(r) == len(hdr) and r[2] == "-" and r[0].isdigit() and lo <= int(r[0]) <= hi:
        d = {hdr[i]: r[i] for i in stall_cols if r[i] not in ("", "0")}
        if d.get("# Samples"):
            top = sorted(((int(v), k) for k, v in d.items() if k.startswith("stall_") and v.isdigit()), reverse=True)[:5]
            print(f"{r[0]:>4} samples {d.get('# Samples','0'):>6} inst {d.get('Instructions Executed','0'):>9}  " +
                  " ".join(f"{k[6:]}={v}" for v, k in top) + "   | " + r[1].strip()[:70])
