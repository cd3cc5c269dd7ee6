#!/usr/bin/env python
"""cuobjdump -sass of lib/libr3dgw_b200.so -> per-kernel opcode histogram (profiles/r02_sass_opcodes.txt).
Evidence for the instruction-level claims in DESIGN.md section 3 (UBLKCP bulk copies, FFMA2 / FMUL2 packed FP32,
REDG...F32x4 vector reductions, bare MUFU.RCP, DFMA in the FP64 geometry chain); runs without a GPU."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "relightable3dgaussians-w_b200", "lib", "libr3dgw_b200.so")
out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_sass_opcodes.txt")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
demangle = lambda n: subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n   # noqa: E731
kern, hist, arch = None, collections.OrderedDict(), set()
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = m.group(1)
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s*arch = (\S+)", line)
    if m:
        arch.add(m.group(1))
    m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        hist[kern][m.group(1)] += 1
KEY = ("UBLKCP", "UTMALDG", "SYNCS", "FFMA2", "FMUL2", "FADD2", "RED", "ATOM", "MUFU.RCP", "MUFU.EX2", "DFMA", "DMUL", "DADD",
       "LDS.128", "LDG.E.128", "STG.E.128", "SHFL", "VOTE", "MATCH", "UTCHMMA", "UTCMMA", "HMMA")
with open(out, "w") as f:
    f.write(f"# cuobjdump -sass {os.path.relpath(lib, ROOT)}; arch {sorted(arch)}; {len(hist)} kernels\n")
    f.write("# per kernel: instruction count, the 12 most frequent opcodes, then the opcodes DESIGN.md cites\n")
    total = collections.Counter()
    for k, h in hist.items():
        total.update(h)
        name = demangle(k)
        name = re.sub(r"\((?:bool|int|unsigned int)\)", "", name)      # template-argument casts: <(bool)1> -> <1>
        name = re.sub(r"\(.*", "", name).replace("void r3dgw::", "").replace("r3dgw::", "")
        n = sum(h.values())
        f.write(f"\n== {name}  [{n} instructions]\n")
        f.write("   top: " + ", ".join(f"{o} {c}" for o, c in h.most_common(12)) + "\n")
        cited = []
        for key in KEY:
            c = sum(v for o, v in h.items() if o.startswith(key) or ("." + key) in o)
            if c:
                cited.append(f"{key} {c}")
        f.write("   cited: " + (", ".join(cited) if cited else "-") + "\n")
    f.write("\n== whole library\n")
    for key in KEY:
        c = sum(v for o, v in total.items() if o.startswith(key) or ("." + key) in o)
        f.write(f"   {key:<10} {c}\n")
print("wrote", out)
