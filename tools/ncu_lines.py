"""Hot source lines of a kernel from an ncu capture taken with --import-source on (binary built with -lineinfo):
    python tools/ncu_lines.py gpurun_out/shade.ncu-rep [capture index] [top n]
Per source file and per line: share of executed warp-instructions and of stall samples."""
import collections, csv, subprocess, sys

rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
agg, samp, text = collections.Counter(), collections.Counter(), {}
cur, hdr, seen, capture = None, None, set(), 0
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        f = r[1].split("/")[-1]
        if f in seen:            # the file list starts over: next capture
            capture += 1
            seen = set()
        seen.add(f)
        cur = f
        continue
    if r[0] == "Line No":
        hdr = r
        iI, iS = hdr.index("Instructions Executed"), hdr.index("# Samples")
        continue
    if r[0] == "Function Name" or hdr is None or capture != which or len(r) <= max(iI, iS) or r[2] != "-":
        continue
    try:
        n, s = int(r[iI]), int(r[iS])
    except ValueError:
        continue
    k = (cur, int(r[0]))
    agg[k] += n
    samp[k] += s
    text[k] = r[1].strip()[:100]
tot, ts = max(1, sum(agg.values())), max(1, sum(samp.values()))
print(f"capture {which}: {tot} warp-instructions, {ts} stall samples\n")
print("| file | % instructions | % stall samples |\n|---|---|---|")
bf, sf = collections.Counter(), collections.Counter()
for k, v in agg.items():
    bf[k[0]] += v
    sf[k[0]] += samp[k]
for f, v in bf.most_common():
    print(f"| {f} | {v / tot * 100:.2f} | {sf[f] / ts * 100:.2f} |")
print("\n| % instructions | % stall samples | line | source |\n|---|---|---|---|")
for k, v in agg.most_common(top):
    print(f"| {v / tot * 100:.2f} | {samp[k] / ts * 100:.2f} | {k[0]}:{k[1]} | `{text[k]}` |")
