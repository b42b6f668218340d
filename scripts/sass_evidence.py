"""SASS evidence of the built library (cuobjdump -sass): per kernel, static counts of the tcgen05 / TMA / mbarrier /
cp.async / multimem / packed-fp32 instructions, and for each kernel family an EXCERPT of the real listing around the first
UTCHMMA / LDTM / UTMALDG / LDGMC site.  Writes profiles/sass_evidence.md."""
import collections, glob, os, re, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = glob.glob(os.path.join(ROOT, "pg*", "lib", "libb200dp.so"))[0]
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
WANT = ["UTCHMMA", "LDTM", "UTCBAR", "UTMALDG", "UTMASTG", "SYNCS", "LDGSTS", "LDGMC", "STGMC", "RED", "ATOMG", "ATOMS", "FFMA2",
        "FADD2", "HMNMX2", "FFMA", "LDS", "STS", "LDG", "STG"]
kern, counts = None, collections.OrderedDict()
listing = collections.OrderedDict()
for line in txt.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = m.group(1)
        counts[kern] = collections.Counter()
        listing[kern] = []
        continue
    if kern and re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
        listing[kern].append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", line).rstrip())
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and kern:
        op = m.group(1)
        for w in WANT:
            if op == w or (w in ("RED", "SYNCS") and op.startswith(w)):
                counts[kern][w] += 1
                break
out = ["# SASS evidence per kernel (`python scripts/sass_evidence.py`: cuobjdump -sass of libb200dp.so, sm_100a)", "",
       "tcgen05.mma -> UTCHMMA, tcgen05.ld -> LDTM, tcgen05.commit -> UTCBAR, TMA load/store -> UTMALDG/UTMASTG, mbarrier -> SYNCS,",
       "cp.async -> LDGSTS, multimem.ld_reduce -> LDGMC.E.ADD.F32x4 (multimem.st is an STG.E.128.STRONG.SYS to the multicast",
       "mapping: the NVSwitch replicates it), red.global -> RED,",
       "fma.rn.f32x2 -> FFMA2, min/max.bf16x2 -> HMNMX2.", "", "## Static instruction counts", ""]
for k, c in counts.items():
    name = re.sub(r"b200dp::(\(anonymous namespace\)::)?", "", demangle(k))
    name = re.sub(r"\(.*", "", name)
    if not c:
        continue
    out.append("- `" + name + "`: " + ", ".join(f"{w} {c[w]}" for w in WANT if c[w]))
# ---- excerpts of the real listing: 4 lines before and 5 after the first site of each marker, one kernel per family
out += ["", "## Listing excerpts (address, instruction) around the first site of each marker", ""]
FAMILIES = [("pw_gemm_kernel", ["UTMALDG", "UTCHMMA", "UTCBAR", "LDTM", "UTMASTG"]), ("pw_wgrad_kernel", ["LDGSTS", "UTCHMMA", "LDTM"]),
            ("dw_tile_bwd_kernel", ["UTMALDG", "FFMA2"]), ("dw_tile_bwd2_kernel", ["UTMALDG", "FFMA2"]),
            ("dw_tile_fwd_kernel", ["UTMALDG", "HMNMX2"]), ("comm_adam_kernel<0, 2, 8>", ["LDGMC", "STG.E.128.STRONG.SYS"])]
for fam, marks in FAMILIES:
    cands = [k for k in listing if fam in demangle(k)]
    if not cands:
        continue
    k = max(cands, key=lambda q: sum(counts[q][m] for m in marks))
    out.append("### `" + re.sub(r"\(.*", "", re.sub(r"b200dp::(\(anonymous namespace\)::)?", "", demangle(k))) + "`")
    for mk in marks:
        idx = [i for i, l in enumerate(listing[k]) if re.search(r"\b" + mk, l)]
        if not idx:
            continue
        i = idx[0]
        out.append("")
        out.append(f"first `{mk}` of {len(idx)}:")
        out.append("```")
        out += [l.strip() for l in listing[k][max(0, i - 4): i + 6]]
        out.append("```")
    out.append("")
open(os.path.join(ROOT, "profiles", "sass_evidence.md"), "w").write("\n".join(out) + "\n")
print("\n".join(l for l in out if any(t in l for t in ("UTCHMMA", "UTMALDG", "FFMA2"))))
