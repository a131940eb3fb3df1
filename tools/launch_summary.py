"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
    python tools/launch_summary.py gpurun_out/launches.csv"""
import collections, csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
agg = collections.OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) < 10:
        continue
    name = re.sub(r"\(.*", "", r[4])[:64]
    v = float(r[-1].replace(",", ""))
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v
tot = sum(t for _, t in agg.values())
for k, (n, t) in agg.items():
    print(f"{k:64s} n={n:4d} total={t / 1e6:9.3f} ms avg={t / n / 1e3:9.1f} us  {100 * t / tot:5.1f}%")
