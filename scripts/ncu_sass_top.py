"""Top SASS instructions by stall samples from an .ncu-rep (source page, sass view)."""
import csv, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name",
                      "regex:" + pat, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
for i, r in enumerate(rows):
    if r and r[0] == 'Address':
        h = r; start = i + 1; break
isrc, isamp, iex = h.index('Source'), h.index('# Samples'), h.index('Instructions Executed')
stall_cols = [k for k in h if k.startswith('stall_') and 'Not Issued' not in k]
data = []
for idx, r in enumerate(rows[start:]):
    try:
        data.append((int(r[isamp]), int(r[iex]), idx, r[isrc][:80], {k: int(r[h.index(k)] or 0) for k in stall_cols}))
    except Exception:
        pass
tot = sum(d[0] for d in data)
print("total samples", tot, "total inst", sum(d[1] for d in data))
for d in sorted(data, reverse=True)[:n]:
    top = sorted(d[4].items(), key=lambda kv: -kv[1])[:2]
    print("%5.1f%% samp ex=%8d  #%d %s  %s" % (100 * d[0] / tot, d[1], d[2], d[3], top))
