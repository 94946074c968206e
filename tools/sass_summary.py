"""Counts of the SASS mnemonics that prove what each kernel of libgcn_b200.so runs on (B200_PROFILING.md):
UTC*MMA (tcgen05.mma), LDTM/STTM (tcgen05.ld/st), UTMALDG (tensor-map TMA), UBLKCP (bulk copy), LDGSTS (cp.async),
FFMA2 (fma.rn.f32x2), SYNCS (mbarrier), plus registers per kernel.   python tools/sass_summary.py > profiles/r02_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "gcn_b200", "lib", "libgcn_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
res = subprocess.run(["cuobjdump", "--dump-resource-usage", lib], capture_output=True, text=True).stdout
regs = {}
for m in re.finditer(r"Function (\S+):\s*\n\s*REG:(\d+) STACK:(\d+) SHARED:(\d+)", res):
    regs[m.group(1)] = (int(m.group(2)), int(m.group(3)), int(m.group(4)))
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
KEYS = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UBLKCP", "LDGSTS", "FFMA2", "FFMA", "SYNCS", "HMMA"]
cur, counts = None, collections.OrderedDict()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"^\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        counts[cur]["_total"] += 1
        for k in KEYS:
            if op == k or op.startswith(k + "."):
                counts[cur][k] += 1
tot = collections.Counter()
print(f"SASS summary of {os.path.relpath(lib, ROOT)} ({len(counts)} kernels); columns: instructions, registers, then non-zero mnemonic counts")
for fn, c in counts.items():
    name = demangle(fn)
    name = re.sub(r"\(.*", "", name)[:110]
    r = regs.get(fn, ("?", "?", "?"))
    hits = " ".join(f"{k}={c[k]}" for k in KEYS if c[k])
    print(f"{c['_total']:6d} instr  {str(r[0]):>3s} regs  {name:110s} {hits}")
    tot.update({k: c[k] for k in KEYS})
print("TOTAL " + " ".join(f"{k}={tot[k]}" for k in KEYS))
