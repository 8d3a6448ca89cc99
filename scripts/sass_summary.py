"""Opcode evidence for the Blackwell-native kernels: per kernel of bist_b200/lib/libbist_b200.so, counts of the SASS mnemonics that
prove tcgen05 (UTC*MMA), TMEM loads (LDTM), TMA (UTMALDG / UTMASTG), the legacy tensor path (HMMA) and the commit / barrier
instructions.  Usage: python scripts/sass_summary.py > profiles/sass_summary.txt   (no GPU needed)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "bist_b200", "lib", "libbist_b200.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
ops = ["UTCHMMA", "UTCBAR", "LDTM", "UTMALDG", "UTMASTG", "UBLKCP", "HMMA", "SYNCS", "LDGSTS", "MUFU.EX2"]
kern, cur = collections.OrderedDict(), None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        kern[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        kern[cur]["_n"] += 1
        for o in ops:
            if m.group(1).startswith(o):
                kern[cur][o] += 1
demangle = subprocess.run(["c++filt"], input="\n".join(kern), capture_output=True, text=True).stdout.splitlines()
print("%-60s %6s " % ("kernel (sm_100a SASS of libbist_b200.so)", "instr") + " ".join("%8s" % o for o in ops))
for (k, c), name in zip(kern.items(), demangle):
    if not any(c[o] for o in ops[:7]):
        continue
    name = name.replace("(anonymous namespace)::", "").replace("bist::", "")
    name = re.sub(r"\(.*", "", name).replace("void ", "")
    print("%-60s %6d " % (name[:60], c["_n"]) + " ".join("%8d" % c[o] for o in ops))
print("\n(kernels without tensor-core / TMA / TMEM instructions are omitted; %d kernels in the library)" % len(kern))
