"""SASS evidence per kernel: counts of the Blackwell-native mnemonics (tcgen05 = UTC*MMA, TMA =
UTMALDG/UTMASTG/UTMAREDG, tcgen05.ld = LDTM, multimem = *MC* / MULTIMEM) in the built library."""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
so = ROOT / "distributed_pytorch_cookbook_b200" / "_C" / "libb200gpt.so"
out = subprocess.run(["cuobjdump", "-sass", str(so)], capture_output=True, text=True).stdout
pats = {"tcgen05.mma.cta_group::2": r"\bUTC[A-Z]*MMA\.2CTA", "TMA load .cta_group::2": r"UTMALDG\.[0-9A-Z]*\.2CTA",
        "commit multicast": r"UTCBAR\.2CTA\.MULTICAST", "tcgen05.st": r"\bSTTM", "tcgen05.mma A-from-TMEM": r"\bUTC[A-Z]*MMA[.A-Z0-9]*\s+tmem\[",
        "tcgen05.mma": r"\bUTC[A-Z]*MMA", "tcgen05.ld": r"\bLDTM", "tcgen05.alloc": r"UTCATOMSWS", "tcgen05.commit": r"UTCBAR",
        "TMA load": r"UTMALDG", "TMA store": r"UTMASTG", "TMA reduce": r"UTMAREDG", "mbarrier": r"SYNCS\.",
        "multimem.ld_reduce": r"LDG[A-Z.]*MC|LDGMC", "multimem.st": r"STG[A-Z.]*MC|STGMC|REDG[A-Z.]*MC",
        "HMMA (legacy)": r"\bHMMA", "cp.async": r"LDGSTS"}
counts = collections.OrderedDict()
name = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(anonymous namespace\)::", "", name).split("(")[0]
        counts[name] = collections.Counter()
        continue
    if name is None:
        continue
    for k, p in pats.items():
        if re.search(p, line):
            counts[name][k] += 1
print(f"SASS evidence for {so.name} (cuobjdump -sass, sm_100a)\n")
for n, c in counts.items():
    if c:
        print(f"{n[:70]:70s} " + "  ".join(f"{k}={v}" for k, v in c.items()))
