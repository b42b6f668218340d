"""ncu launch list with DRAM bytes -> profiles/step_roofline.md: per kernel family and per launch, measured DRAM
traffic, time, achieved bandwidth and its fraction of the measured HBM copy bandwidth (MEASURED_PEAKS.json).
usage: python scripts/step_roofline.py launches.csv out.md [launches_per_step]"""
import collections, csv, json, os, re, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peak = 6553.0
try:
    mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    peak = float(mp.get("hbm_gbs", peak))
except Exception:
    pass
rows = list(csv.reader(open(sys.argv[1], errors="ignore")))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
hdr = rows[hi]
idc, kn, mn, mv, mu = hdr.index("ID"), hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
launch = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) <= mv:
        continue
    d = launch.setdefault(r[idc], {"name": r[kn]})
    v = float(r[mv].replace(",", ""))
    u = r[mu].lower()
    if r[mn].startswith("gpu__time"):
        d["us"] = v / 1e3 if u.startswith("ns") else (v if u.startswith("us") else v * 1e3 if u.startswith("ms") else v * 1e6)
    else:
        mult = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
        d[r[mn]] = v * mult
L = list(launch.values())
per_step = int(sys.argv[3]) if len(sys.argv) > 3 else len(L) // 3
L = L[-per_step:]                                   # the last step
def short(n):
    n = re.sub(r"b200dp::(<unnamed>::|\(anonymous namespace\)::)?", "", n)
    return re.sub(r"^void ", "", re.sub(r"\(.*", "", n))
agg = collections.OrderedDict()
for d in L:
    a = agg.setdefault(short(d["name"]), [0.0, 0.0, 0.0, 0])
    a[0] += d.get("us", 0); a[1] += d.get("dram__bytes_read.sum", 0); a[2] += d.get("dram__bytes_write.sum", 0); a[3] += 1
tot_us = sum(a[0] for a in agg.values()); tot_b = sum(a[1] + a[2] for a in agg.values())
out = [f"# DRAM traffic and achieved bandwidth of one native training step (batch 128, 1x B200)", "",
       f"`ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none` over "
       f"`scripts/one_step.py` (graphs off, last of three steps; ncu serialises launches and starts each with a cold L2, so "
       f"times are upper bounds of the in-graph times).  Roofline denominator: {peak:.0f} GB/s measured HBM copy bandwidth "
       f"(`MEASURED_PEAKS.json`).", "",
       f"**Step total: {len(L)} launches, {tot_us:.0f} us, {tot_b / 1e9:.2f} GB of DRAM traffic "
       f"-> {tot_b / 1e9 / peak * 1e6:.0f} us at the copy bandwidth = {100 * tot_b / 1e9 / peak * 1e6 / tot_us:.0f} % of the step "
       f"(the same step inside its CUDA graph takes less than the ncu sum: see the bench line).**", "",
       "| kernel | launches | us | DRAM read MB | DRAM write MB | GB/s | % of HBM copy bw |", "|---|---|---|---|---|---|---|"]
for name, (us, rd, wr, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    bw = (rd + wr) / 1e9 / (us * 1e-6) if us > 0 else 0
    out.append(f"| `{name}` | {n} | {us:.1f} | {rd / 1e6:.1f} | {wr / 1e6:.1f} | {bw:.0f} | {100 * bw / peak:.0f} % |")
out += ["", "## Every launch, in execution order", "", "| # | kernel | us | read MB | write MB | GB/s | % |", "|---|---|---|---|---|---|---|"]
for i, d in enumerate(L):
    us, rd, wr = d.get("us", 0), d.get("dram__bytes_read.sum", 0), d.get("dram__bytes_write.sum", 0)
    bw = (rd + wr) / 1e9 / (us * 1e-6) if us > 0 else 0
    out.append(f"| {i} | `{short(d['name'])}` | {us:.1f} | {rd / 1e6:.1f} | {wr / 1e6:.1f} | {bw:.0f} | {100 * bw / peak:.0f} % |")
open(sys.argv[2], "w").write("\n".join(out) + "\n")
print("\n".join(out[4:24]))
