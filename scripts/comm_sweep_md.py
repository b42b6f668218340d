"""Raw output of tests/mp_comm_check.py (bytes,algo,fused_us,nccl_us) -> markdown with bus bandwidth and roofline fractions.
usage: python scripts/comm_sweep_md.py raw.txt W out.md
Roofline of the fused step = max(link time, HBM time): link bytes per GPU and direction / 770 GB/s (NVLink 5, ~0.9 TB/s
nominal per direction, ~770 GB/s achievable payload), HBM bytes / measured copy bandwidth."""
import json, os, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
raw, W, out = sys.argv[1], int(sys.argv[2]), sys.argv[3]
hbm = 6547.5
try:
    hbm = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
LINK = 770.0
rows = []
for line in open(raw):
    p = line.strip().split(",")
    if len(p) == 4 and p[0].isdigit():
        rows.append((int(p[0]), p[1], float(p[2]), float(p[3])))
def model(nbytes, algo):
    f = (W - 1) / W
    if algo == "one_shot":      # every rank pulls all peers' gradients, updates everything locally
        link = (W - 1) * nbytes
        mem = nbytes * (1 + 4 + 3 + 1)          # own grads read by me (+ served to peers from L2/HBM), p/m/v read, p/m/v written, zero
    elif algo == "two_shot":    # pull W-1 slices, push the updated slice to W-1 peers
        link = 2 * f * nbytes
        mem = nbytes * (1 + 1 + 1) + 6 * nbytes / W     # grads served, weights received, zero; slice p/m/v read + written
    else:                       # nvls: reduced slice in + others' weights in ; my grads out (pulled by the switch) + my slice out once
        link = nbytes * max(1 / W + f, f + 1 / W)
        mem = nbytes * (1 + 1 + 1) + 6 * nbytes / W
    return link, mem
md = [f"# Fused all-reduce + Adam step vs NCCL all-reduce alone, {W}x B200 (device-timed, max over ranks, 20 calls back to back)", "",
      "`python -m b200dp.launch -np %d -- tests/mp_comm_check.py` (the sweep at its end).  fused = ONE kernel: flag barrier, "
      "all-reduce of fp32 gradients, 1/W, Adam on fp32 master weights + moments, second barrier, gradient zeroing.  "
      "NCCL = `dist.all_reduce` of the same fp32 vector only (no optimiser, no zeroing).  8,970,496 B is the MobileNetV2 "
      "arena of this framework (the shipped configuration).  Roofline = max(link bytes per GPU and direction / 770 GB/s, "
      f"HBM bytes / {hbm:.0f} GB/s); at small sizes the step is the latency of two cross-GPU barriers and a launch, not bandwidth." % W,
      "", "| bytes | algo | fused us | NCCL us | fused / NCCL | bus GB/s (2(W-1)/W x bytes / t) | link bytes MB | roofline us | fraction of roofline |",
      "|---|---|---|---|---|---|---|---|---|"]
for nbytes, algo, fu, nc in rows:
    link, mem = model(nbytes, algo)
    roof = max(link / LINK / 1e3, mem / hbm / 1e3)      # us
    bus = 2 * (W - 1) / W * nbytes / (fu * 1e-6) / 1e9
    md.append(f"| {nbytes:,} | {algo} | {fu:.1f} | {nc:.1f} | {fu / nc:.2f} | {bus:.0f} | {link / 1e6:.1f} | {roof:.1f} | {100 * roof / fu:.0f} % |")
open(out, "w").write("\n".join(md) + "\n")
print("\n".join(md[4:]))
