"""Aggregate an `ncu --page source --csv --print-source sass` dump: executed warp-instructions and stall samples
per opcode, and per contiguous address region (loops show up as plateaus of equal execution counts)."""
import csv, sys, collections, re
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
iS, iE, iSamp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
ops = collections.Counter(); samp = collections.Counter()
recs = []
for r in rows[2:]:
    if len(r) <= iE or r[0] in ("Kernel Name", "Address"): 
        if r and r[0] == "Kernel Name": break          # first launch only
        continue
    src = r[iS].strip()
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_]+)", src)
    op = m.group(2) if m else src.split()[0]
    n = int(r[iE] or 0); s = int(r[iSamp] or 0)
    ops[op] += n; samp[op] += s
    recs.append((src, op, n, s))
tot = sum(ops.values()); ts = sum(samp.values())
print(f"total warp-instructions {tot}, samples {ts}")
for op, n in ops.most_common(28):
    print(f"{op:12s} {n:12d} {100*n/tot:5.1f}%   samples {100*samp[op]/max(ts,1):5.1f}%")
if len(sys.argv) > 2:
    # regions: group consecutive instructions with similar exec count
    print("\nregions (start idx, len, exec count each, share of total, share of samples):")
    i = 0
    while i < len(recs):
        j = i; n0 = recs[i][2]
        while j < len(recs) and (abs(recs[j][2] - n0) <= 0.02 * max(n0, 1)): j += 1
        n = sum(x[2] for x in recs[i:j]); s = sum(x[3] for x in recs[i:j])
        if n > 0.005 * tot:
            mix = collections.Counter(x[1] for x in recs[i:j])
            print(f"{i:5d} {j-i:5d} {n0:10d} {100*n/tot:5.1f}% {100*s/max(ts,1):5.1f}%  " + " ".join(f"{k}:{v}" for k, v in mix.most_common(10)))
        i = j
