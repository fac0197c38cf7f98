"""Counts of the tensor-core / TMA / TMEM / async-copy / atomic / fence mnemonics per kernel of libdie.so (sm_100a).

    python tools/sass_summary.py > profiles/sass_r02.txt
"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed-inference-engine_b200", "die", "_lib", "libdie.so")
PAT = re.compile(r"\b(UTCHMMA|UTCQMMA|UTCMMA|UTCBAR|UTCATOMSWS|LDTM(?:\.\w+)*|UTMALDG(?:\.\w+)*|UTMASTG(?:\.\w+)*|UBLKCP(?:\.\w+)*|"
                 r"LDGSTS(?:\.\w+)*|ATOMG\.E\.CAS\.128|SYNCS\.\w+(?:\.\w+)*|MEMBAR\.\w+\.\w+|CCTL\.IVALL)\b")
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
fn = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
names = iter(fn)
print("# cuobjdump -sass die/_lib/libdie.so (sm_100a): counts of the tensor-core / TMA / TMEM / async-copy / atomic mnemonics per kernel")
print("# UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UTMALDG = cp.async.bulk.tensor (TMA), LDGSTS = cp.async")
cur, counts = None, collections.Counter()


def flush():
    if cur and counts:
        print(cur[:110])
        print("    " + "  ".join(f"{k} x{v}" for k, v in sorted(counts.items())))


for line in sass.split("\n"):
    if "Function :" in line:
        flush()
        cur, counts = next(names), collections.Counter()
        continue
    m = PAT.search(line)
    if m:
        counts[m.group(1)] += 1
flush()
