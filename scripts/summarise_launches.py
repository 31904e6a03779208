"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv): per-kernel totals of the LAST step
(everything after the last random_block_kernel launch, which opens a solve).
    python scripts/summarise_launches.py gpurun_out/launches.csv [--all]"""
import collections, csv, re, sys

rows, hdr = [], None
for r in csv.reader(open(sys.argv[1], errors="replace")):
    if "Kernel Name" in r:
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        if d.get("Metric Name") == "gpu__time_duration.sum":
            v = float(d["Metric Value"].replace(",", ""))
            v *= {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}[d["Metric Unit"]]
            rows.append((re.sub(r"\(.*", "", d["Kernel Name"]).replace("n2v2r::", "").replace("void ", ""), v))
start = 0
if "--all" not in sys.argv:
    for i, (k, _) in enumerate(rows):
        if k.startswith("random_block_kernel"):
            start = i
agg = collections.OrderedDict()
for k, v in rows[start:]:
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(v[1] for v in agg.values())
print(f"{len(rows) - start} launches, {tot:.3f} ms of GPU time (launches {start}..{len(rows) - 1} of the capture)")
print(f"{'kernel':44s} {'launches':>8s} {'total ms':>10s} {'avg us':>9s} {'share':>7s}")
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{k[:44]:44s} {v[0]:8d} {v[1]:10.3f} {1e3 * v[1] / v[0]:9.1f} {100 * v[1] / tot:6.1f}%")
