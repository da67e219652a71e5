"""Per-kernel SASS mnemonic counts of libtsp.so (cuobjdump -sass): which kernels carry the TMA /
bulk-copy / mbarrier / packed-FMA instructions and that none carries a tensor-core instruction.
usage: python tools/sass_summary.py [libtsp.so] > profiles/r02_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "torch_sputnik_b200", "lib", "libtsp.so")
WATCH = ("UTMALDG", "UBLKCP", "SYNCS", "FFMA2", "FHFMA", "FFMA", "LDS", "LDG", "STG", "SHFL", "LDGSTS", "MUFU", "HMMA", "UTCMMA", "IMMA", "LDTM")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
dem = {}
counts = collections.OrderedDict()
cur = None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        op = m.group(1)
        counts[cur]["_all"] += 1
        for w in WATCH:
            if op == w:
                counts[cur][w] += 1
names = subprocess.run(["c++filt"], input="\n".join(counts), capture_output=True, text=True).stdout.splitlines()
groups = collections.OrderedDict()
for mangled, name in zip(counts, names):
    short = re.sub(r"^void ", "", name)
    short = re.sub(r"^tsp::", "", short)
    family = re.sub(r"<.*", "", re.sub(r"\(.*", "", short))
    g = groups.setdefault(family, {"variants": 0, "c": collections.Counter()})
    g["variants"] += 1
    g["c"].update(counts[mangled])
print(f"SASS summary of {os.path.relpath(lib, ROOT)} (sm_100a), `cuobjdump -sass`, summed over the template instantiations of each kernel")
print(f"{'kernel':34s} {'inst':>5s} {'instr':>8s} " + " ".join(f"{w:>7s}" for w in WATCH))
tot = collections.Counter()
for fam, g in groups.items():
    tot.update(g["c"])
    print(f"{fam[:34]:34s} {g['variants']:5d} {g['c']['_all']:8d} " + " ".join(f"{g['c'][w]:7d}" for w in WATCH))
print(f"{'TOTAL':34s} {sum(g['variants'] for g in groups.values()):5d} {tot['_all']:8d} " + " ".join(f"{tot[w]:7d}" for w in WATCH))
print("\nUTMALDG = cp.async.bulk.tensor (TMA tensor-map load), UBLKCP = cp.async.bulk (descriptor-free bulk copy), SYNCS = mbarrier,")
print("FFMA2 = fma.rn.f32x2, FHFMA = fma.rn.f32.f16 / .bf16 (16-bit factors, fp32 addend and result).  HMMA / UTCMMA / IMMA / LDTM (tensor-core / TMEM) must be 0: north_star excludes tensor cores for this path.")
