"""SASS opcode census of libb200lora.so: the opcodes that prove the Blackwell-native paths, per kernel.
    python tools/sass_census.py > profiles/r02_sass_census.md"""
import collections
import re
import subprocess
import sys

SO = "diffusion_finetuning_b200/libb200lora.so"
KEEP = re.compile(r"^(UTCHMMA|UTCBAR|UTCATOMSWS|LDTM|STTM|UTMALDG|UTMASTG|UTMACCTL|UBLKCP|SYNCS|ACQBULK|ELECT|HMMA|LDSM|F2FP|FFMA2|"
                  r"HADD2\.BF16|HMUL2\.BF16|MEMBAR\.ALL\.GPU|CCTL\.IVALL)")
out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
per = collections.OrderedDict()
cur = None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"^void ", "", name)
        name = re.sub(r"<.*", "", name)
        name = re.sub(r"\(.*", "", name)
        cur = per.setdefault(name, collections.Counter())
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Za-z0-9_.]+)", line)
    if m and cur is not None and KEEP.match(m.group(1)):
        cur[m.group(1)] += 1
print("# r02 — SASS opcode census of `libb200lora.so` (sm_100a)\n")
print("`python tools/sass_census.py` (`cuobjdump -sass`), opcodes that prove the Blackwell-native paths, counted per kernel (all template\n"
      "instantiations of one kernel summed).  UTCHMMA = tcgen05.mma (`.2CTA` = cta_group::2), LDTM = tcgen05.ld (TMEM -> registers),\n"
      "UTMALDG / UTMASTG = TMA tensor load / store (`.IM2COL` = im2col mode), UBLKCP = cp.async.bulk (1-D bulk copies: the merge ring, the\n"
      "pair kernel's relay), UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, HMMA / LDSM = mma.sync / ldmatrix (the HBM-bound token\n"
      "reductions), F2FP = packed 16-bit conversions, FFMA2 = fma.rn.f32x2 (merge), MEMBAR.ALL.GPU / CCTL.IVALL = what cluster-scope\n"
      "release / acquire compiles to (only reachable with B200LORA_CLUSTER_SCOPE=1; the default relay mode does not execute them).\n")
print("| kernel | opcode counts |")
print("|---|---|")
tot = collections.Counter()
for k, c in per.items():
    if not c:
        continue
    tot.update(c)
    print(f"| `{k}` | " + ", ".join(f"{op} x{n}" for op, n in c.most_common()) + " |")
print("\nTotals: " + ", ".join(f"{op} x{n}" for op, n in tot.most_common()))
