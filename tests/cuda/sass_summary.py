"""Opcode histogram per kernel of the shipped library (for profiles/sass_summary.txt):
    python tests/cuda/sass_summary.py > profiles/sass_summary.txt
Shows which kernels carry tcgen05 (UTCHMMA / UTCBAR), TMEM (LDTM / STTM), bulk copies (UBLKCP) and mbarriers (SYNCS)."""
import collections
import os
import re
import subprocess
import sys

LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tts_b200", "csrc", "libtts_b200.so")
KEY = ["UTCHMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "STTM", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "HMMA", "FFMA",
       "HFMA2", "MUFU", "LDG", "STG", "LDS", "STS", "LDL", "STL", "BAR", "ACQBULK", "NANOSLEEP"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = kernels.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m and cur is not None:
            cur[m.group(1)] += 1
    demangle = subprocess.run(["cu++filt"] + list(kernels), capture_output=True, text=True).stdout.splitlines()
    print("SASS opcode counts per kernel of tts_b200/csrc/libtts_b200.so (cuobjdump -sass; static instruction counts)")
    print("columns: " + " ".join(KEY) + " | total")
    tot = collections.Counter()
    for (name, ops), nice in zip(kernels.items(), demangle):
        nice = re.sub(r"\(anonymous namespace\)::|<unnamed>::|ttsb::|\((int|bool)\)", "", nice)
        nice = re.sub(r"\(.*$", "", nice).replace("void ", "")
        print(f"{nice[:70]:<70} " + " ".join(f"{ops.get(k, 0):>5}" for k in KEY) + f" | {sum(ops.values())}")
        tot.update(ops)
    print(f"{'ALL':<70} " + " ".join(f"{tot.get(k, 0):>5}" for k in KEY) + f" | {sum(tot.values())}")


if __name__ == "__main__":
    main()
