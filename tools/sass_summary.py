"""Per-kernel SASS evidence: counts of the Blackwell-native mnemonics in every kernel of libpeftb200.so.

    python tools/sass_summary.py > profiles/rNN_sass_summary.txt

UTCHMMA = tcgen05.mma (.2CTA = cta_group::2), LDTM / STTM = tcgen05.ld / st (TMEM), UTMALDG / UTMASTG / UTMAREDG = TMA loads /
stores / reduce-adds, UTCBAR = tcgen05.commit, FFMA2 / FMUL2 / FADD2 = packed fp32 math, HMMA = legacy mma.sync.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "parameter-efficient-fine-tuning-for-vision-transformers_b200", "csrc", "libpeftb200.so")
KEYS = ["UTCHMMA.2CTA", "UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UTCBAR", "ELECT", "R2UR", "FFMA2", "FMUL2", "FADD2",
        "MUFU", "HMMA", "ATOMS", "RED"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    counts = collections.OrderedDict()
    name = None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(.*", "", name)
            counts[name] = collections.Counter()
            continue
        if name is None:
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if not m:
            continue
        op = m.group(1)
        c = counts[name]
        c["total"] += 1
        for k in KEYS:
            if op == k or op.startswith(k + "."):
                c[k] += 1
    print("# SASS mnemonic counts per kernel of libpeftb200.so (cuobjdump -sass); UTCHMMA includes the .2CTA forms")
    print(f"{'kernel':100s} {'instr':>6s} " + " ".join(f"{k[:9]:>9s}" for k in KEYS))
    tot = collections.Counter()
    for name, c in counts.items():
        tot.update(c)
        if c["total"] == 0:
            continue
        print(f"{name[:100]:100s} {c['total']:6d} " + " ".join(f"{c[k]:9d}" for k in KEYS))
    print(f"{'TOTAL':100s} {tot['total']:6d} " + " ".join(f"{tot[k]:9d}" for k in KEYS))


if __name__ == "__main__":
    main()
