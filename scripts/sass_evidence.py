"""Per-kernel SASS mnemonic counts of the built library (cuobjdump -sass): which kernels really contain tcgen05 / TMA /
mbarrier / cp.async / packed-fp32 instructions.  Writes profiles/sass_evidence.md."""
import collections, glob, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = glob.glob(os.path.join(ROOT, "pg*", "lib", "libb200dp.so"))[0]
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
WANT = ["UTCHMMA", "LDTM", "UTCBAR", "UTMALDG", "UTMASTG", "SYNCS", "LDGSTS", "RED", "ATOMG", "ATOMS", "FFMA2", "FADD2",
        "HMNMX2", "FFMA", "LDS", "STS", "LDG", "STG"]
kern, counts = None, collections.OrderedDict()
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = m.group(1)
        counts[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and kern:
        op = m.group(1)
        for w in WANT:
            if op == w or (w in ("RED", "SYNCS") and op.startswith(w)):
                counts[kern][w] += 1
                break
out = ["# SASS evidence per kernel (`python scripts/sass_evidence.py`: cuobjdump -sass of libb200dp.so, sm_100a)", "",
       "tcgen05.mma -> UTCHMMA, tcgen05.ld -> LDTM, tcgen05.commit -> UTCBAR, TMA load/store -> UTMALDG/UTMASTG, mbarrier -> SYNCS,",
       "cp.async -> LDGSTS, red.global -> RED, fma.rn.f32x2 -> FFMA2, min/max.bf16x2 -> HMNMX2.  Static instruction counts.", ""]
for k, c in counts.items():
    name = re.sub(r"b200dp::(\(anonymous namespace\)::)?", "", demangle(k))
    name = re.sub(r"\(.*", "", name)
    if not c:
        continue
    out.append("- `" + name + "`: " + ", ".join(f"{w} {c[w]}" for w in WANT if c[w]))
open(os.path.join(ROOT, "profiles", "sass_evidence.md"), "w").write("\n".join(out) + "\n")
print("\n".join(l for l in out if any(t in l for t in ("UTCHMMA", "UTMALDG", "FFMA2"))))
