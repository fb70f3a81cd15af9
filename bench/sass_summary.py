"""Static SASS evidence for every kernel of libdryad_kernels.so: registers, instruction count and
the hardware-path mnemonics that matter (tcgen05 = UTCHMMA/UTCBAR/LDTM, TMA = UTMALDG/UBLKCP/UTMAREDG,
bulk copies, multimem = LDGMC, reductions = REDG/ATOMG, 256-bit loads, FMNMX3 ...).
Usage: python bench/sass_summary.py > profiles/sass_summary.txt   (needs only cuobjdump)."""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "dryad_b200/_lib/libdryad_kernels.so"
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
res = subprocess.run(["cuobjdump", "-res-usage", lib], capture_output=True, text=True).stdout
regs = {}
cur = None
for line in res.splitlines():
    m = re.search(r"Function (\S+):", line)
    if m:
        cur = m.group(1)
    m = re.search(r"REG:(\d+).*?SHARED:(\d+)", line)
    if m and cur:
        regs[cur] = (int(m.group(1)), int(m.group(2)))
KEY = ["UTCHMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "UBLKRED", "LDGMC", "REDG", "ATOMG", "ATOMS",
       "SYNCS", "LDG.E.ENL2.256", "LDG.E.128", "STG.E.128", "FMNMX3", "MATCH", "SHFL", "BAR.SYNC", "LDS.128", "ERRBAR", "MEMBAR", "CCTL"]
funcs = collections.OrderedDict()
name = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = m.group(1)
        funcs[name] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and name:
        funcs[name][m.group(1)] += 1
dem = subprocess.run(["cu++filt"] + list(funcs), capture_output=True, text=True).stdout.splitlines() if funcs else []
print(f"# SASS summary of {lib} (sm_100a): {len(funcs)} kernels\n")
for (mangled, cnt), pretty in zip(funcs.items(), dem or funcs):
    total = sum(cnt.values())
    head = pretty.split("(")[0]
    short = head if "<" not in head or ">" in head else pretty[:pretty.index(">") + 1]
    r = regs.get(mangled, ("?", "?"))
    keys = []
    for k in KEY:
        n = sum(v for op, v in cnt.items() if op.startswith(k))
        if n:
            keys.append(f"{k}={n}")
    print(f"{short}\n    regs={r[0]} static_smem={r[1]} instructions={total}  " + " ".join(keys))
