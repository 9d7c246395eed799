"""Counts the SASS mnemonics that prove which hardware path each kernel of libsplitnn.so uses
(B200_PROFILING.md: UTCHMMA = tcgen05.mma, UTMALDG = TMA tensor load, LDTM/STTM = tensor-memory
load/store, UTCBAR = tcgen05.commit, FFMA2 = packed fp32 FMA, LDGSTS = cp.async, SYNCS = mbarrier).

    python profiles/sass_summary.py > profiles/sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "splitlearning_b200", "libsplitnn.so")
OPS = ["UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "STTM", "UTCBAR", "SYNCS", "LDGSTS", "FFMA2", "FFMA", "HFMA2", "MUFU", "SHFL",
       "LDS", "STS", "LDG", "STG", "ATOM", "RED", "BAR", "NANOSLEEP"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    demangle = {}
    names = sorted(set(re.findall(r"Function : (\S+)", sass)))
    out = subprocess.run(["c++filt"] + names, capture_output=True, text=True).stdout.splitlines()
    for n, d in zip(names, out):
        demangle[n] = re.sub(r"\(.*", "", d).replace("sn::", "")
    counts, size, cur = collections.defaultdict(collections.Counter), collections.Counter(), None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = demangle[m.group(1)]
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            size[cur] += 16
            op = m.group(1)
            for o in OPS:
                if op == o or op.startswith(o + "."):
                    counts[cur][o] += 1
                    break
    print("# cuobjdump -sass splitlearning_b200/libsplitnn.so, static instruction counts per kernel (sm_100a)")
    print("kernel,code_KB," + ",".join(OPS))
    for k in sorted(size, key=lambda k: -size[k]):
        print(f"\"{k}\",{size[k] / 1024:.0f}," + ",".join(str(counts[k][o]) for o in OPS))


if __name__ == "__main__":
    sys.exit(main())
