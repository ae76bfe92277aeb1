"""Opcode histogram of the sm_100a SASS in libgenseg_b200.so (`cuobjdump -sass`), per kernel: the evidence that the
tensor-core kernels use tcgen05 (UTCHMMA / UTCHMMA.2CTA, UTCBAR, TMEM loads LDTM), TMA (UTMALDG / UTMASTG) and that no
floating-point atomics (RED / ATOM .ADD.F32) are left.  Writes profiles/sass_summary.txt."""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "semantic_segmentation_b200/libgenseg_b200.so"
out = sys.argv[2] if len(sys.argv) > 2 else "profiles/sass_summary.txt"
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", txt)), capture_output=True, text=True).stdout.split("\n")
kern = None
hist = collections.OrderedDict()
idx = -1
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        idx += 1
        kern = re.sub(r"\(.*", "", names[idx].replace("(anonymous namespace)::", "")).replace("void gsb::", "")
        hist.setdefault(kern, collections.Counter())
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        hist[kern][m.group(1)] += 1
KEY = ("UTCHMMA", "UTCQMMA", "UTCBAR", "UTCCP", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UTMACCTL", "SYNCS", "UBLKCP",
       "HMMA", "RED", "ATOM", "ATOMG", "FFMA2", "FMUL2", "FADD2", "DFMA", "DADD", "MUFU")
with open(out, "w") as f:
    f.write("# cuobjdump -sass %s : per-kernel instruction count and the opcodes that identify the hardware path\n" % lib)
    f.write("# (tcgen05.mma = UTCHMMA[.2CTA], tcgen05.commit = UTCBAR, tcgen05.ld = LDTM, TMA = UTMALDG / UTMASTG, mbarrier = SYNCS)\n")
    tot = collections.Counter()
    for k, h in hist.items():
        n = sum(h.values())
        sel = {}
        for op, c in h.items():
            base = op.split(".")[0]
            if base in KEY:
                name = op if base in ("UTCHMMA", "UTCBAR", "RED", "ATOM", "ATOMG", "UTMALDG", "UTMASTG") else base
                sel[name] = sel.get(name, 0) + c
                tot[name] += c
        f.write("%-90s %6d  %s\n" % (k[:90], n, " ".join("%s=%d" % kv for kv in sorted(sel.items()))))
    f.write("\n# totals over the library\n")
    for op, c in sorted(tot.items()):
        f.write("%-40s %d\n" % (op, c))
    fa = [op for op in tot if op.split(".")[0] in ("RED", "ATOM", "ATOMG") and (".F32" in op or ".F64" in op or ".F16" in op)]
    f.write("\n# floating-point atomics: %s\n" % (", ".join(fa) if fa else "none"))
print(open(out).read()[-1200:])
