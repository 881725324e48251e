"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel.
usage: python profiles/summarize_launches.py gpurun_out/launches.csv > profiles/<round>_launches_summary.txt
Per-launch times under ncu are cold-cache and serialised: compare SHARES with bench.py's own
`roofline.share_of_step`, not absolutes."""
import collections
import csv
import re
import sys


def main(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    tot, cnt = collections.defaultdict(float), collections.Counter()
    for row in csv.DictReader(lines):
        name = re.sub(r"\(.*", "", row["Kernel Name"])
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}.get(row["Metric Unit"], 1.0)
        tot[name] += v
        cnt[name] += 1
    T = sum(tot.values())
    print(f"# {path}: {sum(cnt.values())} launches, {T / 1e6:.3f} ms total device time under ncu")
    print(f"{'total ms':>10} {'share':>7} {'n':>6} {'avg us':>10}  kernel")
    for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
        print(f"{v / 1e6:10.3f} {100 * v / T:6.2f}% {cnt[k]:6d} {v / cnt[k] / 1e3:10.1f}  {k[:110]}")


if __name__ == "__main__":
    main(sys.argv[1])
