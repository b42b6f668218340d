"""ncu --csv launch list (gpu__time_duration.sum per kernel launch) -> markdown table by kernel.
usage: python scripts/summarise_launches.py launches.csv out.md [skip_fraction]   (skip_fraction: leading warm-up part)"""
import collections, csv, re, sys

rows = list(csv.reader(open(sys.argv[1], errors="ignore")))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
hdr = rows[hi]
kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
data = []
for r in rows[hi + 1:]:
    if len(r) <= mv:
        continue
    v = float(r[mv].replace(",", ""))
    unit = r[mu]
    us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
    data.append((r[kn], us))
skip = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
data = data[int(len(data) * skip):]
agg = collections.OrderedDict()
for name, us in data:
    name = re.sub(r"b200dp::(<unnamed>::|\(anonymous namespace\)::)?", "", name)
    name = re.sub(r"^void ", "", re.sub(r"\(.*", "", name))
    a = agg.setdefault(name, [0.0, 0])
    a[0] += us
    a[1] += 1
total = sum(a[0] for a in agg.values())
out = [f"{len(data)} launches, {total:.0f} us in total.", "", "| kernel | us | share | launches |", "|---|---|---|---|"]
for name, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    out.append(f"| `{name}` | {us:.1f} | {100 * us / total:.1f}% | {n} |")
open(sys.argv[2], "w").write("\n".join(out) + "\n")
print("\n".join(out[:14]))
