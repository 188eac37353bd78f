"""Regenerates profiles/sass_counts_r2.txt: per-kernel counts of the SASS mnemonics that show what the shipped
libctdet_b200.so runs on (cuobjdump -sass; no GPU needed):

  UTCHMMA / UTCQMMA   tcgen05.mma (5th-gen tensor cores)        LDTM / STTM        tcgen05.ld / st (TMEM)
  UTCBAR              tcgen05.commit -> mbarrier                UBLKCP             cp.async.bulk (TMA 1-D bulk copy)
  UTMALDG / UTMASTG   cp.async.bulk.tensor (tensor-map TMA)     SYNCS              mbarrier operations
  REDG / ATOMG        global reductions / atomics               ACQBULK / PREEXIT  griddepcontrol.wait / launch_dependents
  HMMA / IMMA         legacy mma.sync (must be 0)               MUFU               special-function unit

    python scripts/sass_counts.py
"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "detectron2_centernet_b200", "libctdet_b200.so")
OUT = os.path.join(ROOT, "profiles", "sass_counts_r2.txt")
WANT = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "REDG", "ATOMG", "ATOMS",
        "ACQBULK", "PREEXIT", "HMMA", "IMMA", "MUFU", "LDG", "STG", "LDS", "STS", "BAR"]
sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
arch = sorted(set(re.findall(r"arch = (sm_\w+)", sass)))
per, total, name = collections.OrderedDict(), collections.Counter(), None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        per[name] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and name:
        op = m.group(1)
        per[name]["_instructions"] += 1
        for w in WANT:
            if op == w or op.startswith(w + "."):
                per[name][w] += 1
                total[w] += 1
with open(OUT, "w") as f:
    f.write("SASS mnemonic counts of detectron2_centernet_b200/libctdet_b200.so (cuobjdump -sass), arch = %s\n" % ", ".join(arch))
    f.write("regenerate: python scripts/sass_counts.py\n\n")
    f.write("whole library: " + "  ".join("%s=%d" % (w, total[w]) for w in WANT) + "\n\n")
    for k, c in per.items():
        f.write("%-60s instrs=%-6d " % (k[:60], c["_instructions"]) + "  ".join("%s=%d" % (w, c[w]) for w in WANT if c[w]) + "\n")
print(open(OUT).read()[:3000])
