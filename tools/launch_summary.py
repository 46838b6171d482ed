"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list (profiles/*_bench_launches.csv)."""
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 5]
hdr = next(r for r in rows if "Kernel Name" in r)
iN, iV, iM, iU = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Name"), hdr.index("Metric Unit")
tot, cnt, mx = collections.Counter(), collections.Counter(), collections.Counter()
for r in rows:
    if r is hdr or r[iM] != "gpu__time_duration.sum":
        continue
    v = float(r[iV].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[iU], 1e-3)
    tot[r[iN]] += v; cnt[r[iN]] += 1; mx[r[iN]] = max(mx[r[iN]], v)
T = sum(tot.values())
print(f"(cold-cache, serialised launch times; {sum(cnt.values())} launches; total {T:.1f} us)")
print(" count     total us     avg us     max us  share  kernel")
for k, v in tot.most_common():
    print(f"{cnt[k]:6d} {v:12.1f} {v / cnt[k]:10.1f} {mx[k]:10.1f} {100 * v / T:5.1f}%  {k[:110]}")
