"""Summarise an `ncu -i report.ncu-rep --page raw --csv` export (optionally gzipped): one line per captured launch
with the metrics the roofline discussion uses.
usage: python profiles/summarize_ncu_raw.py gpurun_out/raw.csv.gz [label ...] > profiles/<round>_ncu_<what>.txt
Labels (optional) name the launches in capture order."""
import csv
import gzip
import io
import re
import sys

COLS = [
    ("time us", "gpu__time_duration.sum"),
    ("tensor %", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"),
    ("sm thr %", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("dram rd MB", "dram__bytes_read.sum"),
    ("dram wr MB", "dram__bytes_write.sum"),
    ("dram %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("smem wavefr %", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"),
    ("issue act %", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    ("stall no_inst", "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio"),
    ("stall long_sb", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
    ("regs", "launch__registers_per_thread"),
    ("grid", "launch__grid_size"),
    ("block", "launch__block_size"),
]
SCALE = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}


def main(path, labels):
    f = gzip.open(path, "rt") if path.endswith(".gz") else open(path)
    rows = list(csv.reader(f))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = []
    for _, m in COLS:
        idx.append(hdr.index(m) if m in hdr else None)
    print("# " + path)
    print(f"{'#':>3} {'kernel':38s} " + " ".join(f"{n:>13s}" for n, _ in COLS) + "  label")
    for n, r in enumerate(data):
        name = re.sub(r"\(.*", "", r[4]).replace("void <unnamed>::", "")[:38]
        vals = []
        for i in idx:
            if i is None or r[i] == "":
                vals.append("-")
                continue
            v = float(r[i].replace(",", "")) * SCALE.get(units[i], 1.0)
            vals.append(f"{v:.1f}" if abs(v) < 1e5 else f"{v:.3g}")
        print(f"{n:3d} {name:38s} " + " ".join(f"{v:>13s}" for v in vals) + "  " + (labels[n] if n < len(labels) else ""))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2:])
