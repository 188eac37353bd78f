"""Executed-instruction mix by SASS opcode from an .ncu-rep (source page, sass view)."""
import csv, subprocess, sys, collections
rep, pat = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name",
                      "regex:" + pat, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
for i, r in enumerate(rows):
    if r and r[0] == 'Address':
        h = r; start = i + 1; break
isrc, iex = h.index('Source'), h.index('Instructions Executed')
mix = collections.Counter()
for r in rows[start:]:
    try:
        ex = int(r[iex])
    except Exception:
        continue
    toks = r[isrc].split()
    op = toks[1] if toks and toks[0].startswith('@') else (toks[0] if toks else '?')
    mix[op.split('.')[0] + ('.' + op.split('.')[1] if op.startswith('MUFU') else '')] += ex
tot = sum(mix.values())
print('total warp instructions', tot)
for k, v in mix.most_common(int(sys.argv[3]) if len(sys.argv) > 3 else 30):
    print('%-14s %10d %5.1f%%' % (k, v, 100 * v / tot))
