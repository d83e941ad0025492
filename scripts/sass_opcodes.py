"""Per-kernel counts of the SASS opcodes that prove the Blackwell path (tcgen05 = UTCHMMA / UTC*MMA, TMA = UTMALDG /
UTMASTG, TMEM = LDTM / STTM) and of HMMA (mma.sync), from the built library.
usage: python scripts/sass_opcodes.py > profiles/sass_opcodes.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "copilot-for-consensus_b200", "libb200_retrieval.so")
OPS = ["UTCHMMA", "UTCHMMA.2CTA", "UTMALDG", "UTMASTG", "LDTM", "STTM", "UTCBAR", "HMMA", "MUFU.EX2"]


def main() -> None:
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    counts: dict[str, collections.Counter] = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = re.sub(r"\(.*", "", name).replace("b200::", "").replace("void ", "")
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        for op in OPS:
            if re.search(r"\b" + re.escape(op) + r"\b", line):
                counts[cur][op] += 1
    print("SASS opcode counts per kernel of copilot-for-consensus_b200/libb200_retrieval.so (cuobjdump -sass, sm_100a)")
    print("%-58s" % "kernel" + "".join("%14s" % o for o in OPS))
    tot = collections.Counter()
    for k, c in counts.items():
        print("%-58s" % k[:57] + "".join("%14d" % c[o] for o in OPS))
        tot.update(c)
    print("%-58s" % "TOTAL" + "".join("%14d" % tot[o] for o in OPS))


if __name__ == "__main__":
    main()
