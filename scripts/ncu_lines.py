"""Per-CUDA-source-line executed-instruction / stall-sample shares from an .ncu-rep (cuda,sass view)."""
import csv, subprocess, sys, collections
rep, pat = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                      "regex:" + pat, "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
cur_file = None
hdr = None
agg = collections.OrderedDict()
for r in csv.reader(out.splitlines()):
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; ie = hdr.index("Instructions Executed"); isamp = hdr.index("# Samples"); continue
    if hdr is None or r[0] in ("Function Name",): continue
    if r[2] != "-": continue            # SASS rows carry an address; CUDA rows have '-'
    try: e = int(r[ie]); s = int(r[isamp])
    except ValueError: continue
    k = (cur_file, r[0], r[1].strip()[:100])
    a = agg.setdefault(k, [0, 0]); a[0] += e; a[1] += s
tot = sum(a[0] for a in agg.values()) or 1; tots = sum(a[1] for a in agg.values()) or 1
print("total warp-instr", tot, "samples", tots)
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][int(__import__("os").environ.get("SORTCOL","0"))])[:int(sys.argv[4]) if len(sys.argv) > 4 else 40]:
    print("%5.1f%% inst %5.1f%% samp  %s:%s  %s" % (100 * a[0] / tot, 100 * a[1] / tots, k[0], k[1], k[2]))
