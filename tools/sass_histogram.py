"""Per-kernel SASS opcode histogram of libvc_b200.so (VERDICT r1 "weak" #11): which Blackwell-only instructions each kernel
uses — UTCHMMA (tcgen05.mma; `.2CTA` = cta_group::2), LDTM/STTM (tcgen05.ld/st), UTMALDG (TMA tensor load), UBLKCP (TMA
bulk copy), UTCBAR (tcgen05.commit), SYNCS (mbarrier), HMMA (legacy mma.sync: must be 0).
    python tools/sass_histogram.py > profiles/sass_opcodes_r02.txt"""
import collections
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(HERE, "versecrafter_b200", "libvc_b200.so")
KEYS = ["UTCHMMA", "UTCHMMA.2CTA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "SYNCS", "HMMA", "MUFU.EX2", "STG", "LDG"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    archs = sorted(set(re.findall(r"arch = (sm_\w+)", sass)))
    per = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = re.sub(r"\(.*", "", cur.replace("(anonymous namespace)::", "")).replace("void ", "")
            while cur in per:
                cur += "'"
            per[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            op = m.group(1)
            base = op.split(".")[0]
            per[cur][base] += 1
            if op.startswith("UTCHMMA") and ".2CTA" in op:
                per[cur]["UTCHMMA.2CTA"] += 1
            if op.startswith("MUFU.EX2"):
                per[cur]["MUFU.EX2"] += 1
            per[cur]["_total"] += 1
    print(f"# {os.path.relpath(LIB, HERE)}: code objects {archs}; {len(per)} kernels; counts are static SASS instructions")
    print("# " + " ".join(f"{k:>12}" for k in ["total"] + KEYS) + "  kernel")
    tot = collections.Counter()
    for name, c in per.items():
        print("  " + " ".join(f"{c.get(k, 0):>12}" for k in ["_total"] + KEYS) + "  " + name)
        tot.update(c)
    print("# " + " ".join(f"{tot.get(k, 0):>12}" for k in ["_total"] + KEYS) + "  ALL")
    assert archs == ["sm_100a"], archs
    assert tot["HMMA"] == 0


if __name__ == "__main__":
    sys.exit(main())
