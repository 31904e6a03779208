"""Per-kernel counts of the SASS mnemonics that prove tcgen05 / TMA / TMEM use in libn2v2r_b200.so (no GPU needed).
    python scripts/sass_summary.py > profiles/r02_sass_summary.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "node2vec2rank_b200", "libn2v2r_b200.so")
sass = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
KEYS = ["UTCHMMA", "UTCQMMA", "UTMALDG", "UBLKCP", "LDTM", "UTCBAR", "UTCATOMSWS", "SYNCS", "FENCE.VIEW.ASYNC", "UCGABAR", "STS", "LDS", "BAR.SYNC"]
cur, counts, variants = None, collections.OrderedDict(), collections.defaultdict(collections.Counter)
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if not m:
        continue
    op = m.group(1)
    for k in KEYS:
        if op.startswith(k):
            counts[cur][k] += 1
    if op.startswith("UTCHMMA") or op.startswith("UTCQMMA"):
        v = "+".join(sorted(set(re.findall(r"A_KEEP|A_REUSE", line)))) or "plain"
        variants[cur][("2CTA " if ".2CTA" in line else "") + v] += 1
print("kernel".ljust(64), " ".join(k.rjust(9) for k in KEYS))
for name, c in counts.items():
    if any(c[k] for k in KEYS[:6]):
        print(name[:64].ljust(64), " ".join(str(c[k]).rjust(9) for k in KEYS))
print("\nMMA variants (A-operand collector flags):")
for name, c in variants.items():
    print(" ", name[:64].ljust(64), dict(c))
