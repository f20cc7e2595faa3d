"""Instruction evidence for the Blackwell-native claims: per kernel of libmakani_b200.so, the counts of the SASS
mnemonics that only tcgen05 / TMEM / TMA code produces (B200_PROFILING.md: tcgen05.mma -> UTC*MMA, tcgen05.ld -> LDTM,
cp.async.bulk.tensor -> UTMALDG / UTMASTG, cp.async.bulk -> UBLKCP) and of the legacy tensor path (HMMA: must be 0).

    python tools/sass_summary.py > profiles/sass_summary.txt        (no GPU needed: cuobjdump reads the .so)"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "makani_b200", "libmakani_b200.so")
WANT = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTMAPF", "SYNCS", "HMMA", "LDGSTS", "FFMA2", "FADD2", "FMUL2"]

sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
counts = collections.OrderedDict()
cur = None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur).replace("void ", "").replace("sfno::", "")
        counts[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1).split(".")[0]
        counts[cur]["_all"] += 1
        for w in WANT:
            if op.startswith(w):
                counts[cur][w] += 1

print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)}  (sm_100a)   - tools/sass_summary.py")
print("# kernel | instructions | " + " | ".join(WANT))
tot = collections.Counter()
for k, c in counts.items():
    if not any(c[w] for w in WANT):
        continue
    print(f"{k[:110]} | {c['_all']} | " + " | ".join(str(c[w]) for w in WANT))
    tot.update(c)
print("TOTAL (kernels listed) | " + str(tot["_all"]) + " | " + " | ".join(str(tot[w]) for w in WANT))
others = [k for k, c in counts.items() if not any(c[w] for w in WANT)]
print(f"# {len(others)} further kernels use none of these (plain LDG/STG bandwidth kernels): " + ", ".join(sorted(set(o[:40] for o in others)))[:1500])
