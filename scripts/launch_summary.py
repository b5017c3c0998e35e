"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list: python scripts/launch_summary.py list.csv"""
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = next(r for r in rows if "Kernel Name" in r)
ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows:
    if r is hdr or r[ik] == "Kernel Name":
        continue
    try:
        v = float(r[iv].replace(",", ""))
    except ValueError:
        continue
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3}.get(r[iu], 1.0)
    k = r[ik].split("(")[0][:110]
    tot[k] += v
    cnt[k] += 1
all_us = sum(tot.values())
print(f"{sum(cnt.values())} launches, {all_us / 1e3:.1f} ms of kernel time (cold-cache, serialised: compare shares)")
for k, v in tot.most_common():
    print(f"{k:112s} {cnt[k]:6d} {v:12.1f} us {100 * v / all_us:6.2f}%")
