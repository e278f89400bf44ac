"""Per-kernel histogram of the SASS mnemonics that prove tcgen05 / TMEM / TMA (B200_PROFILING.md): UTCHMMA / UTCIMMA (tcgen05.mma
kind::f16 / kind::i8), LDTM / STTM (tcgen05.ld / st), UTMALDG (TMA tensor load), UBLKCP (bulk copy), SYNCS (mbarrier), REDG /
RED (fire-and-forget reductions), from `cuobjdump -sass` of the shipped library.
    python tools/sass_opcodes.py > profiles/r02_sass_opcodes.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "umx.cpp_b200", "libumx_cuda.so")
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
KEYS = ["UTCHMMA", "UTCIMMA", "UTCQMMA", "LDTM", "STTM", "UTCBAR", "UTMALDG", "UTMASTG", "UBLKCP", "UBLKPF", "SYNCS", "RED", "ATOM",
        "HMMA", "IMMA", "FFMA", "MUFU", "BAR", "LDS", "STS", "LDG", "STG", "LDL", "STL"]
cur, counts, total = None, collections.OrderedDict(), collections.Counter()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().replace("(anonymous namespace)::", "").split("(")[0]
        counts[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1).split(".")[0]
        total[cur] += 1
        for k in KEYS:
            if op == k or (k in ("RED", "ATOM") and op.startswith(k)):
                counts[cur][k] += 1
print(f"SASS opcode counts per kernel, sm_100a, {os.path.basename(LIB)} ({os.path.getsize(LIB)} bytes)\n")
print(f"{'kernel':58s} {'instr':>7s}  " + " ".join(f"{k:>7s}" for k in KEYS))
for name, c in counts.items():
    print(f"{name[:58]:58s} {total[name]:7d}  " + " ".join(f"{c.get(k, 0):7d}" for k in KEYS))
