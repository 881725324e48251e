"""Counts of the Blackwell-native SASS mnemonics per kernel of the in-tree library (cuobjdump -sass):
UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG = TMA loads, UTCBAR = tcgen05.commit.
usage: python profiles/sass_counts.py > profiles/<round>_sass_counts.txt"""
import collections, os, re, subprocess, sys
so = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "speakingstyle_b200", "lib", "libspeakingstyle_b200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
pat = re.compile(r"\b(UTC[A-Z]*MMA(?:\.2CTA)?|LDTM|STTM|UTMALDG|UTCBAR|HMMA|SYNCS\.PHASECHK|FFMA)\b")
cur, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", name.replace("(anonymous namespace)::", "").replace("void ", ""))
        counts[cur] = collections.Counter()
        continue
    if cur:
        for mm in pat.findall(line):
            counts[cur][mm] += 1
cols = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTCBAR", "HMMA", "FFMA"]
print(f"{'kernel':70s} " + " ".join(f"{c:>12s}" for c in cols))
tot = collections.Counter()
for k, c in counts.items():
    if not any(c[x] for x in cols[:6]):
        continue
    print(f"{k[:70]:70s} " + " ".join(f"{c[x]:12d}" for x in cols))
    tot.update(c)
print(f"{'TOTAL (kernels with tcgen05 / TMA instructions)':70s} " + " ".join(f"{tot[x]:12d}" for x in cols))
